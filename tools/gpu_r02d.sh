#!/bin/bash
# round 2, 2-GPU pass: NCCL parity tests through the C ABI, bench at N=2 (strong record with the gather timed), PCIe ceiling
set -u
mkdir -p gpurun_out
nvidia-smi -L
nvidia-smi topo -m > gpurun_out/r02d_topo.txt 2>&1
timeout 600 python -m pytest tests/test_partition_gpu.py -m gpu -q > gpurun_out/r02d_pytest_partition.log 2>&1; echo "partition rc=$?"; tail -8 gpurun_out/r02d_pytest_partition.log | cut -c1-300
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > gpurun_out/r02d_bench_n2.json 2> gpurun_out/r02d_bench_n2.err; echo "bench2 rc=$?"; tail -c 1500 gpurun_out/r02d_bench_n2.err; python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r02d_bench_n2.json').read().strip().splitlines()[-1]); print(json.dumps(d.get('strong'),indent=1)[:3500]); print(d['value'], d['e2e'])
except Exception as e: print('no n2 line', e)
PY
nvcc -O2 -std=c++17 -o /tmp/pcie_peak tools/pcie_peak.cu -lpthread && timeout 300 /tmp/pcie_peak > gpurun_out/r02d_pcie_peak.json 2> gpurun_out/r02d_pcie_peak.err; echo "pcie rc=$?"; head -c 3000 gpurun_out/r02d_pcie_peak.json
