#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 200 ncu --set full --clock-control none --import-source on -k regex:mass3d_kron -c 1 -s 2 -o gpurun_out/r02z_kron3_p3 -f python tools/prof_new_kernels.py kron3 > /dev/null 2>&1; echo "kron3 rc=$?"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:mass2d_kron -c 1 -s 2 -o gpurun_out/r02z_gradmirror_p16 -f python tools/prof_new_kernels.py gradmirror > /dev/null 2>&1; echo "grad rc=$?"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:l2proj_fd2d -c 1 -s 2 -o gpurun_out/r02z_fd_c4 -f python tools/prof_new_kernels.py fd_c4 > /dev/null 2>&1; echo "fd rc=$?"
timeout 200 ncu --set full --clock-control none --import-source on -k regex:l2proj_fd2d -c 1 -s 4 -o gpurun_out/r02z_fd_c5 -f python tools/bench_fd.py --only c5 --fd-only > /dev/null 2>&1; echo "fd c5 rc=$?"
