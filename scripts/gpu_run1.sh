#!/usr/bin/env bash
# First GPU pass: parity tests, smoke, bench (both arms), launch list + full ncu capture of the sort pass.
set -uo pipefail
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
lscpu | head -20 > gpurun_out/lscpu.txt
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/pytest_gpu.log 2>&1
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/smoke.log 2>&1
( timeout 600 python bench.py ) > gpurun_out/bench.json 2> gpurun_out/bench.err
( timeout 300 python bench.py --impl reference --steps 3 --warmup 1 ) > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
CMD="python bench.py --steps 3 --warmup 3 --no-extras --no-e2e --no-cpu"
( timeout 300 $CMD > gpurun_out/plain.log 2>&1 ) && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_sort.csv $CMD > gpurun_out/ncu1.log 2>&1
( timeout 300 $CMD > gpurun_out/plain2.log 2>&1 ) && \
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:sort_pass -s 24 -c 3 -o gpurun_out/prof_sort_pass $CMD > gpurun_out/ncu2.log 2>&1
echo finished
