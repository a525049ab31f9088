#!/usr/bin/env bash
set -uo pipefail
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -30 ) > gpurun_out/pytest_gpu.log 2>&1
cat gpurun_out/pytest_gpu.log
for wl in q1_fused groupby_c4; do
  ( timeout 600 python bench.py --workload $wl --no-extras --no-cpu --no-e2e --steps 5 ) > gpurun_out/b_$wl.json 2> gpurun_out/b_$wl.err
  tail -2 gpurun_out/b_$wl.err
  python - $wl <<'PY'
import json,sys
try:
    d=json.load(open('gpurun_out/b_%s.json'%sys.argv[1]))
    print(sys.argv[1], "ms/step %.3f"%d["ms_per_step"], "rows/s %.3g"%d["value"], "frac", d["roofline"] and round(d["roofline"]["frac"],3), "kernels", {k:round(v,3) for k,v in d["kernel_ms_per_step"].items()})
except Exception as e: print(sys.argv[1], "failed", e)
PY
done
CMD="python bench.py --workload groupby_c4 --rows 200000000 --no-extras --no-cpu --no-e2e --steps 1 --warmup 3"
( timeout 300 $CMD > gpurun_out/c4_plain.log 2>&1 ) && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gb_hp_agg -s 3 -c 1 -o gpurun_out/prof_agg $CMD > gpurun_out/agg_ncu.log 2>&1
echo "ncu agg rc $?"
CMD="python bench.py --workload q1_fused --no-extras --no-cpu --no-e2e --steps 1 --warmup 3"
( timeout 300 $CMD > gpurun_out/q1_plain.log 2>&1 ) && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gb_fused8 -s 3 -c 1 -o gpurun_out/prof_q1 $CMD > gpurun_out/q1_ncu.log 2>&1
echo "ncu q1 rc $?"
