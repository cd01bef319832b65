#!/bin/bash
# round 2: Kronecker mass kernel (C3 single launch, C4) + fast-diagonalisation projection: parity, bench lines, ncu
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_kform_gpu.py tests/test_interp_gpu.py -m gpu -x -q > gpurun_out/r02g_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02g_pytest.log | cut -c1-250
for w in c3 c4; do
  timeout 600 python bench.py --workload $w --no-cpu-baseline --steps 10 > gpurun_out/r02g_bench_$w.json 2> gpurun_out/r02g_bench_$w.err; echo "bench $w rc=$?"; tail -3 gpurun_out/r02g_bench_$w.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/r02g_bench_$w.json').read().strip().splitlines()[-1]); r=d['roofline']
print('$w', d['value'], d['ms_per_step'], 'frac', r['frac'], 'step_frac', r['step_frac'], 'e2e', d['e2e']['value'] if d.get('e2e') else None)
PY
done
IB200_NO_KRON=1 timeout 600 python bench.py --workload c4 --no-cpu-baseline --no-e2e --steps 10 > gpurun_out/r02g_bench_c4_nokron.json 2>/dev/null; python -c "
import json; d=json.loads(open('gpurun_out/r02g_bench_c4_nokron.json').read().strip().splitlines()[-1]); print('c4 nokron', d['value'], d['roofline']['step_frac'])"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mass2d_kron -c 1 -s 3 -o gpurun_out/r02g_kron_c3 -f python bench.py --workload c3 --no-cpu-baseline --no-e2e --steps 3 --warmup 3 > gpurun_out/r02g_ncu_c3.log 2>&1; echo "ncu c3 rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mass2d_kron -c 1 -s 3 -o gpurun_out/r02g_kron_c4 -f python bench.py --workload c4 --no-cpu-baseline --no-e2e --steps 3 --warmup 3 > gpurun_out/r02g_ncu_c4.log 2>&1; echo "ncu c4 rc=$?"
