#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02j_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r02j_pytest_gpu.log | cut -c1-250
timeout 900 python bench.py > gpurun_out/r02j_bench_n1.json 2> gpurun_out/r02j_bench_n1.err; echo "bench1 rc=$?"; tail -c 800 gpurun_out/r02j_bench_n1.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02j_bench_n1.json').read().strip().splitlines()[-1])
print('c2', d['value'], d['roofline']['frac'], 'e2e', d['e2e']['value'])
m=d['mass']; print('mass', m['value'], m['roofline']['frac'], 'e2e', m['e2e']['value'], m['e2e']['ms_per_step'])
s=d['strong']; s.pop('config'); print(json.dumps(s, indent=1))
PY
