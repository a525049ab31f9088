#!/usr/bin/env bash
set -uo pipefail
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests/test_full_size_gpu.py -m gpu -q -k "config4" -s 2>&1 | tail -30 ) > gpurun_out/pytest_full.log 2>&1
cat gpurun_out/pytest_full.log
CMD="python bench.py --workload groupby_c4 --rows 200000000 --no-extras --no-cpu --no-e2e --steps 1 --warmup 3"
( timeout 300 $CMD > gpurun_out/c4_plain.log 2>&1 ) && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gb_hp_agg -s 3 -c 1 -o gpurun_out/prof_agg2 $CMD > gpurun_out/agg_ncu.log 2>&1
echo "ncu agg rc $?"
