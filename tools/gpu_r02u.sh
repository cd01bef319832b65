#!/bin/bash
# round 2, final multi-GPU pass: the default bench command exactly as the driver launches it, at N = $1
set -u
N=$1
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N > gpurun_out/r02u_bench_n$N.json 2> gpurun_out/r02u_bench_n$N.err; echo "bench$N rc=$?"; tail -c 600 gpurun_out/r02u_bench_n$N.err | grep -v "OMP\|\*\*\*"
python - <<PY
import json
d=json.loads(open('gpurun_out/r02u_bench_n$N.json').read().strip().splitlines()[-1])
print('c2', d['value'], 'frac', d['roofline']['frac'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'])
m=d['mass']; print('mass', m['value'], m['roofline']['frac'], 'e2e chain', m['e2e']['value'], m['e2e']['ms_per_step'])
s=d['strong']; print('strong', s['value'], 'step', s['ms_per_step'], 'compute', s['compute_only_ms'], 'coll', s['collective_only_ms'], s['collective']['allgather_gbs_per_gpu'], 'single', s['single_gpu_ms'], 'speedup', s['speedup_vs_single_gpu'], 'parity', s['collective_parity'], s['config']['gather_path'][:30])
PY
