#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_parity_gpu.py tests/test_interp_gpu.py tests/test_dropin_gpu.py -m gpu -q > gpurun_out/r02o_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r02o_pytest.log | cut -c1-300
for w in c1 c4; do
timeout 600 python bench.py --workload $w --steps 20 > gpurun_out/r02o_bench_$w.json 2> gpurun_out/r02o_bench_$w.err; echo "bench $w rc=$?"; tail -3 gpurun_out/r02o_bench_$w.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02o_bench_$w.json').read().strip().splitlines()[-1]); r=d['roofline']
print('$w', d['value'], 'ms', d['ms_per_step'], 'frac', r['frac'], 'step', r['step_frac'], 'launches', d['gpu_launches'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], 'cpu', d['cpu_baseline']['value'] if d.get('cpu_baseline') else None)
PY
done
