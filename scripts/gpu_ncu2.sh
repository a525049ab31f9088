#!/usr/bin/env bash
# ncu --set full captures of the two dominant kernels (sort pass, group-by fast path)
set -uo pipefail
mkdir -p gpurun_out
CMD="python bench.py --workload sort_i64 --no-extras --no-cpu --no-e2e --steps 2 --warmup 3"
( timeout 300 $CMD > gpurun_out/sort_plain.log 2>&1 ) && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:sort_pass -s 26 -c 2 -o gpurun_out/prof_sort $CMD > gpurun_out/sort_ncu.log 2>&1
echo "sort ncu rc $?"
CMD="python bench.py --workload groupby_c3 --no-extras --no-cpu --no-e2e --steps 2 --warmup 3"
( timeout 300 $CMD > gpurun_out/gb_plain.log 2>&1 ) && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:gb_bin32 -s 3 -c 1 -o gpurun_out/prof_gb $CMD > gpurun_out/gb_ncu.log 2>&1
echo "gb ncu rc $?"
CMD="python bench.py --workload filter --no-extras --no-cpu --no-e2e --steps 2 --warmup 3"
( timeout 300 $CMD > gpurun_out/filter_plain.log 2>&1 ) && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:filter_apply -s 3 -c 1 -o gpurun_out/prof_filter $CMD > gpurun_out/filter_ncu.log 2>&1
echo "filter ncu rc $?"
