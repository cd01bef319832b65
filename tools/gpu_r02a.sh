#!/bin/bash
# round 2, first GPU pass (2-GPU box): NCCL parity tests through the C ABI, whole -m gpu suite, bench at N=1 and N=2, PCIe ceiling
set -u
mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests/test_partition_gpu.py -m gpu -x -q > gpurun_out/r02a_pytest_partition.log 2>&1; echo "partition rc=$?"; tail -5 gpurun_out/r02a_pytest_partition.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02a_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r02a_pytest_gpu.log
timeout 900 python bench.py > gpurun_out/r02a_bench_n1.json 2> gpurun_out/r02a_bench_n1.err; echo "bench1 rc=$?"; tail -c 1500 gpurun_out/r02a_bench_n1.err; head -c 6000 gpurun_out/r02a_bench_n1.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > gpurun_out/r02a_bench_n2.json 2> gpurun_out/r02a_bench_n2.err; echo "bench2 rc=$?"; tail -c 1500 gpurun_out/r02a_bench_n2.err; python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r02a_bench_n2.json').read().strip().splitlines()[-1]); print(json.dumps(d.get('strong'),indent=1)[:3000]); print(d['value'], d['e2e'])
except Exception as e: print('no n2 line', e)
PY
nvcc -O2 -std=c++17 -o tools/pcie_peak tools/pcie_peak.cu -lpthread && timeout 300 tools/pcie_peak > gpurun_out/r02a_pcie_peak.json 2> gpurun_out/r02a_pcie_peak.err; echo "pcie rc=$?"; head -c 3000 gpurun_out/r02a_pcie_peak.json
