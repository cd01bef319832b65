#!/bin/bash
# round 2: fast-diagonalisation projection -- parity, timings, ncu capture of the C5 kernel
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "l2_project" > gpurun_out/r02f_pytest_fd.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02f_pytest_fd.log | cut -c1-250
timeout 600 python tools/bench_fd.py > gpurun_out/r02f_bench_fd.jsonl 2> gpurun_out/r02f_bench_fd.err; echo "bench_fd rc=$?"; cat gpurun_out/r02f_bench_fd.jsonl; tail -5 gpurun_out/r02f_bench_fd.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:l2proj_fd2d -c 1 -s 4 -o gpurun_out/r02f_fd_c5 -f python tools/bench_fd.py --only c5 --fd-only > gpurun_out/r02f_ncu.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/r02f_ncu.log
