#!/bin/bash
# Runs on the GPU box (under gpurun): parity tests of the "next" rows, their device-resident timings, then the ncu launch
# list and one full capture per kernel (N1 spd_solve, N2 points, N3 dof_derivative).  Outputs land in gpurun_out/.
set -u
mkdir -p gpurun_out
if [ -z "${SKIP_TESTS:-}" ]; then
  timeout 900 python -m pytest tests -m gpu -x -q -k "${TEST_FILTER:-incidence or points or spd or kform}" 2>&1 | tail -3
fi
python tools/bench_next.py | tee gpurun_out/bench_next.jsonl
if [ -n "${NCU_KERNELS:-}" ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_next.csv python tools/bench_next.py > gpurun_out/ncu_list_next.log 2>&1
  for k in $NCU_KERNELS; do
    ncu --set full --clock-control none --import-source on -k regex:"$k" -s 2 -c 1 -f -o gpurun_out/prof_next_$k python tools/bench_next.py > gpurun_out/ncu_full_next_$k.log 2>&1
    tail -1 gpurun_out/ncu_full_next_$k.log | cut -c1-200
  done
fi
ls gpurun_out
