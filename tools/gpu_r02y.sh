#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_partition_gpu.py -m gpu -q > gpurun_out/r02y_pytest_partition.log 2>&1; echo "partition rc=$?"; tail -3 gpurun_out/r02y_pytest_partition.log
bash tools/gpu_r02u.sh 2
cp gpurun_out/r02u_bench_n2.json gpurun_out/r02y_bench_n2.json
