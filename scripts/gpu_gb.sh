#!/usr/bin/env bash
# group-by iteration: parity tests, C3 bench line, ncu capture of the consume kernel at 200 M rows
set -uo pipefail
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_groupby_gpu.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -5 ) > gpurun_out/gb_pytest.log 2>&1
( timeout 300 python bench.py --workload groupby_c3 --no-extras --no-cpu --no-e2e --steps 5 ) > gpurun_out/gb_bench.json 2> gpurun_out/gb_bench.err
CMD="python bench.py --workload groupby_c3 --rows 200000000 --no-extras --no-cpu --no-e2e --steps 3"
if [ "${1:-}" = "ncu" ]; then
( timeout 300 $CMD > gpurun_out/gb_plain.log 2>&1 ) && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:gb_ -s 3 -c 1 -o gpurun_out/prof_gb $CMD > gpurun_out/gb_ncu.log 2>&1
fi
cat gpurun_out/gb_pytest.log; python - <<'PY'
import json
d=json.load(open('gpurun_out/gb_bench.json'))
print("ms/step", d["ms_per_step"], "rows/s %.3g"%d["value"], "roofline", d["roofline"]["achieved"], d["roofline"]["frac"], d["kernel_ms_per_step"])
PY
