#!/bin/bash
# round 2, 1-GPU pass: whole -m gpu suite (incl. drop-in acceptance), default bench line, ncu launch list of the bench
set -u
mkdir -p gpurun_out
nvidia-smi -L
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02c_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/r02c_pytest_gpu.log | cut -c1-300
cat gpurun_out/dropin_package.json gpurun_out/dropin_extension.json 2>/dev/null
timeout 900 python bench.py > gpurun_out/r02c_bench_n1.json 2> gpurun_out/r02c_bench_n1.err; echo "bench1 rc=$?"; tail -c 1500 gpurun_out/r02c_bench_n1.err; head -c 9000 gpurun_out/r02c_bench_n1.json
