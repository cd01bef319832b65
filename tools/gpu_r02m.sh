#!/bin/bash
set -u
mkdir -p gpurun_out
for mode in p2p nccl; do
  if [ $mode = nccl ]; then export IB200_NO_P2P_GATHER=1; fi
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --no-cpu-baseline --no-e2e --no-mass > gpurun_out/r02m_bench_n8_$mode.json 2> gpurun_out/r02m_bench_n8_$mode.err; echo "bench8 $mode rc=$?"; tail -c 400 gpurun_out/r02m_bench_n8_$mode.err | grep -v "OMP\|\*\*\*"
  python - <<PY
import json
d=json.loads(open('gpurun_out/r02m_bench_n8_$mode.json').read().strip().splitlines()[-1])
s=d['strong']; print('$mode', 'value', s['value'], 'step', s['ms_per_step'], 'compute', s['compute_only_ms'], 'coll', s['collective_only_ms'], 'GB/s', s['collective']['allgather_gbs_per_gpu'], 'single', s['single_gpu_ms'], 'speedup', s['speedup_vs_single_gpu'], 'eff', s['efficiency'], 'linkeff', s['link_bound_efficiency'], 'parity', s['collective_parity'])
PY
done
