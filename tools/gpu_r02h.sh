#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_kform_gpu.py tests/test_interp_gpu.py -m gpu -q > gpurun_out/r02h_pytest.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/r02h_pytest.log | cut -c1-250
for w in c4 c3; do
  timeout 600 python bench.py --workload $w --no-cpu-baseline --no-e2e --steps 10 > gpurun_out/r02h_bench_$w.json 2> gpurun_out/r02h_bench_$w.err; echo "bench $w rc=$?"; tail -3 gpurun_out/r02h_bench_$w.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/r02h_bench_$w.json').read().strip().splitlines()[-1]); r=d['roofline']
print('$w', d['value'], d['ms_per_step'], 'frac', r['frac'], 'step_frac', r['step_frac'], 'kernel_ms', r['kernel_ms'])
PY
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mass2d_kron -c 1 -s 3 -o gpurun_out/r02h_kron_c4 -f python bench.py --workload c4 --no-cpu-baseline --no-e2e --steps 3 --warmup 3 > gpurun_out/r02h_ncu_c4.log 2>&1; echo "ncu c4 rc=$?"
