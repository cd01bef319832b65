#!/bin/bash
# round 2 verification pass: whole -m gpu suite, smoke(), default bench, "next rows" timings
set -u
mkdir -p gpurun_out
timeout 420 python -m pytest tests -m gpu -q > gpurun_out/r02x_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r02x_pytest_gpu.log | cut -c1-250
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 400 python bench.py > gpurun_out/r02x_bench_n1.json 2> gpurun_out/r02x_bench_n1.err; echo "bench1 rc=$?"; tail -c 500 gpurun_out/r02x_bench_n1.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02x_bench_n1.json').read().strip().splitlines()[-1])
print('c2', d['value'], d['roofline']['frac'], 'e2e', d['e2e']['value'], 'cpu', d['cpu_baseline']['value'])
m=d['mass']; print('mass', m['value'], m['roofline']['frac'], 'e2e', m['e2e']['value'], 'cpu', m['cpu_baseline']['value'])
s=d['strong']; print('strong', s['value'], s['ms_per_step'], s['roofline']['frac'], s['collective_parity'])
PY
timeout 300 python tools/bench_next.py > gpurun_out/r02x_bench_next.jsonl 2> gpurun_out/r02x_bench_next.err; echo "next rc=$?"; cut -c1-200 gpurun_out/r02x_bench_next.jsonl
timeout 200 python tools/bench_fd.py > gpurun_out/r02x_bench_fd.jsonl 2>/dev/null; echo "fd rc=$?"; timeout 200 python tools/bench_gradmass.py > gpurun_out/r02x_bench_gradmass.jsonl 2>/dev/null; timeout 200 python tools/bench_3d.py > gpurun_out/r02x_bench_3d.jsonl 2>/dev/null; timeout 100 python tools/kron_sweep.py > gpurun_out/r02x_kron_sweep.jsonl 2>/dev/null; echo tools done
