#!/usr/bin/env bash
# round 2: all GPU parity tests, sort + C4 bench lines, ncu --set full of the partition / aggregate / sort-pass kernels
set -uo pipefail
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/pytest_gpu.log 2>&1
cat gpurun_out/pytest_gpu.log
for wl in sort_i64 sort_f64 groupby_c4; do
  ( timeout 600 python bench.py --workload $wl --no-extras --no-cpu --no-e2e --steps 5 ) > gpurun_out/b_$wl.json 2> gpurun_out/b_$wl.err
  python - $wl <<'PY'
import json,sys
try:
    d=json.load(open('gpurun_out/b_%s.json'%sys.argv[1]))
    print(sys.argv[1], "ms/step %.3f"%d["ms_per_step"], "rows/s %.3g"%d["value"], "kernels", {k:round(v,3) for k,v in d["kernel_ms_per_step"].items()})
except Exception as e: print(sys.argv[1], "failed", e)
PY
done
CMD="python bench.py --workload groupby_c4 --rows 200000000 --no-extras --no-cpu --no-e2e --steps 1 --warmup 3"
( timeout 300 $CMD > gpurun_out/c4_plain.log 2>&1 ) && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gb_hp_\(agg\|part\|hist\) -s 15 -c 5 -o gpurun_out/prof_c4 $CMD > gpurun_out/c4_ncu.log 2>&1
echo "ncu c4 rc $?"
CMD="python bench.py --workload sort_i64 --no-extras --no-cpu --no-e2e --steps 1 --warmup 3"
( timeout 300 $CMD > gpurun_out/sort_plain.log 2>&1 ) && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sort_\(pass\|hist\) -s 27 -c 9 -o gpurun_out/prof_sort $CMD > gpurun_out/sort_ncu.log 2>&1
echo "ncu sort rc $?"
ls -la gpurun_out/*.ncu-rep
