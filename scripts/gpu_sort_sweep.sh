#!/usr/bin/env bash
# sort-pass tile-shape sweep: each variant library runs the sort parity tests' core + the C2 bench
set -uo pipefail
mkdir -p gpurun_out
for lib in variants/lib_*.so; do
  name=$(basename $lib .so)
  ( B200_LIB=$PWD/$lib timeout 300 python bench.py --workload sort_i64 --no-extras --no-cpu --no-e2e --steps 5 ) > gpurun_out/sweep_$name.json 2> gpurun_out/sweep_$name.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open('gpurun_out/sweep_%s.json'%n))
    print("%-16s ms/step %.3f  pass %.3f ms  frac %.3f"%(n, d["ms_per_step"], d["roofline"]["avg_launch_ms"], d["roofline"]["frac"]))
except Exception as e: print(n,"failed",e)
PY
done | tee gpurun_out/sort_tile_sweep.txt
