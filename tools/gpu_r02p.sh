#!/bin/bash
# ncu: launch list of the default bench command, full captures of the C4 (staged) Kronecker kernel and of K2-lo in the mass record
set -u
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02p_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r02p_ncu_bench.log 2>&1; echo "ncu list rc=$?"; wc -l gpurun_out/r02p_launches.csv
timeout 600 ncu --set full --clock-control none --import-source on -k regex:mass2d_kron -c 1 -s 3 -o gpurun_out/r02p_kron_c4 -f python bench.py --workload c4 --no-cpu-baseline --no-e2e --steps 3 --warmup 3 > gpurun_out/r02p_ncu_c4.log 2>&1; echo "ncu c4 rc=$?"
