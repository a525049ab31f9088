#!/usr/bin/env bash
# full GPU pass + ncu launch list of the default bench command
set -uo pipefail
bash scripts/gpu_all.sh
( timeout 300 python bench.py --impl reference --steps 2 --warmup 1 ) > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
tail -c 600 gpurun_out/bench_ref.json
CMD="python bench.py --no-extras --no-cpu --no-e2e --steps 3 --warmup 3"
( timeout 300 $CMD > gpurun_out/launch_plain.log 2>&1 ) && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launch.log 2>&1
echo "ncu rc $?"
