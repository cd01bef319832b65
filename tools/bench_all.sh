#!/bin/bash
# One line per workload: value, e2e (run under gpurun).
for w in ${WORKLOADS:-c2 c5 c3 c4}; do
  python bench.py --workload $w --no-cpu-baseline > gpurun_out/all_$w.json 2> gpurun_out/all_$w.err
  python - $w <<'PY'
import json, sys
w = sys.argv[1]
d = json.loads(open(f"gpurun_out/all_{w}.json").read().strip().splitlines()[-1])
r = d["roofline"]
print(w, f"value {d['value']:.4g} ms {d['ms_per_step']:.4f} frac {r['frac']:.3f} step {r['step_frac']:.3f} e2e {d['e2e']['value']:.4g} ({d['e2e']['ms_per_step']:.1f} ms)")
PY
done
