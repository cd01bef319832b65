#!/bin/bash
# round 2, 8-GPU pass: bench at N=8 (strong record with the gather timed), PCIe ceiling with 1/2/4/8 GPUs
set -u
mkdir -p gpurun_out
nvidia-smi -L | wc -l
nvidia-smi topo -m > gpurun_out/r02e_topo.txt 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 > gpurun_out/r02e_bench_n8.json 2> gpurun_out/r02e_bench_n8.err; echo "bench8 rc=$?"; tail -c 1500 gpurun_out/r02e_bench_n8.err; python - <<'PY'
import json
try:
    d=json.loads(open('gpurun_out/r02e_bench_n8.json').read().strip().splitlines()[-1]); s=d.get('strong'); s.pop('config',None); print(json.dumps(s,indent=1)[:3500]); print(d['value'], d['e2e']['value'], d['e2e']['ms_per_step'], d['mass']['value'], d['mass']['e2e']['value'], d['mass']['e2e']['ms_per_step'])
except Exception as e: print('no n8 line', e)
PY
nvcc -O2 -std=c++17 -o /tmp/pcie_peak tools/pcie_peak.cu -lpthread 2>/dev/null && timeout 300 /tmp/pcie_peak > gpurun_out/r02e_pcie_peak.json 2> gpurun_out/r02e_pcie_peak.err; echo "pcie rc=$?"; head -c 4000 gpurun_out/r02e_pcie_peak.json
