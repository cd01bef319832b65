#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py tests/test_partition_gpu.py -m gpu -q -k "fastdiag or partition or nccl or gather" > gpurun_out/r02k_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02k_pytest.log | cut -c1-250
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --no-cpu-baseline > gpurun_out/r02k_bench_n2.json 2> gpurun_out/r02k_bench_n2.err; echo "bench2 rc=$?"; tail -c 600 gpurun_out/r02k_bench_n2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r02k_bench_n2.json').read().strip().splitlines()[-1])
s=d['strong']; s.pop('config'); print(json.dumps(s, indent=1))
PY
