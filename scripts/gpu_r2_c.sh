#!/usr/bin/env bash
# round 2: all GPU parity tests (no -x), bench lines for C4 / Q1 variants
set -uo pipefail
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -q 2>&1 | tail -40 ) > gpurun_out/pytest_gpu.log 2>&1
cat gpurun_out/pytest_gpu.log
for wl in groupby_c4 q1 q1_fused; do
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
