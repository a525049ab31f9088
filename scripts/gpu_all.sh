#!/usr/bin/env bash
# full GPU pass: every parity test, smoke, the default bench and the filter / group-by lines
set -uo pipefail
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5 ) > gpurun_out/pytest_gpu.log 2>&1
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" ) > gpurun_out/smoke.log 2>&1
( timeout 600 python bench.py ) > gpurun_out/bench.json 2> gpurun_out/bench.err
cat gpurun_out/pytest_gpu.log gpurun_out/smoke.log
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print("main", d["config"]["workload"], "ms %.3f"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], "e2e ms", d["e2e"]["ms_per_step"], "cpu", d["cpu_baseline"]["value"])
for k,v in d.get("workloads",{}).items():
    if "error" in v: print(k, v); continue
    print(k, "ms %.3f"%v["ms_per_step"], "rows/s %.3g"%v["rows_per_s"], "frac %.3f"%v["roofline"]["frac"], "kernel ms %.3f"%v["roofline"]["avg_launch_ms"])
PY
