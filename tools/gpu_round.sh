#!/bin/bash
# Runs on the GPU box (under gpurun): bench lines for the workloads named in $BENCH_WORKLOADS, then the ncu launch
# list and one full capture of the dominant kernel for those in $NCU_WORKLOADS.  Outputs land in gpurun_out/.
set -u
mkdir -p gpurun_out
if [ -n "${FP64_PEAK:-}" ]; then
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp64_peak tools/fp64_peak.cu && ./tools/fp64_peak | tee gpurun_out/fp64_peak.json
fi
for w in ${BENCH_WORKLOADS:-c2 c5 c4 c3 c1}; do
  python bench.py --workload $w > gpurun_out/bench_$w.json 2> gpurun_out/bench_$w.err
  tail -c 2200 gpurun_out/bench_$w.json; tail -c 600 gpurun_out/bench_$w.err; echo
done
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
for w in ${NCU_WORKLOADS:-}; do
  $B --workload $w > gpurun_out/plain_$w.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c ${NCU_LIST_COUNT:-80} --csv --log-file gpurun_out/launches_$w.csv $B --workload $w > gpurun_out/ncu_list_$w.log 2>&1
  $B --workload $w > gpurun_out/plain_$w.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:"${NCU_KERNEL:-eval2d|mass2d}" -s ${NCU_SKIP:-3} -c 1 -f -o gpurun_out/prof_$w $B --workload $w > gpurun_out/ncu_full_$w.log 2>&1
  tail -2 gpurun_out/ncu_full_$w.log | cut -c1-300
done
ls gpurun_out
