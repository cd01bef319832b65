#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_partition_gpu.py -m gpu -q -x > gpurun_out/r02l_pytest.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/r02l_pytest.log | cut -c1-300
for mode in p2p nccl; do
  if [ $mode = nccl ]; then export IB200_NO_P2P_GATHER=1; fi
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --no-cpu-baseline --no-e2e --no-mass > gpurun_out/r02l_bench_n2_$mode.json 2> gpurun_out/r02l_bench_n2_$mode.err; echo "bench2 $mode rc=$?"; tail -c 400 gpurun_out/r02l_bench_n2_$mode.err | grep -v OMP
  python - <<PY
import json
d=json.loads(open('gpurun_out/r02l_bench_n2_$mode.json').read().strip().splitlines()[-1])
s=d['strong']; print('$mode', s['config']['gather_path'][:40], 'step', s['ms_per_step'], 'compute', s['compute_only_ms'], 'coll', s['collective_only_ms'], 'GB/s', s['collective']['allgather_gbs_per_gpu'], 'eff', s['efficiency'], 'parity', s['collective_parity'])
PY
done
