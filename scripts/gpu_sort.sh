#!/usr/bin/env bash
# sort iteration: parity tests, C2 bench, optional ncu capture of the pass kernel
set -uo pipefail
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_sort_gpu.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -3 ) > gpurun_out/sort_pytest.log 2>&1
( timeout 300 python bench.py --workload sort_i64 --no-extras --no-cpu --no-e2e --steps 10 ) > gpurun_out/sort_bench.json 2> gpurun_out/sort_bench.err
( timeout 300 python bench.py --workload sort_f64 --no-extras --no-cpu --no-e2e --steps 10 ) > gpurun_out/sort_bench_f64.json 2>> gpurun_out/sort_bench.err
CMD="python bench.py --workload sort_i64 --no-extras --no-cpu --no-e2e --steps 3"
if [ "${1:-}" = "ncu" ]; then
( timeout 300 $CMD > gpurun_out/sort_plain.log 2>&1 ) && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:sort_pass -s 25 -c 2 -o gpurun_out/prof_sort $CMD > gpurun_out/sort_ncu.log 2>&1
fi
cat gpurun_out/sort_pytest.log; python - <<'PY'
import json
for f in ("sort_bench","sort_bench_f64"):
    try:
        d=json.load(open('gpurun_out/%s.json'%f))
        print(f, "ms/step %.3f"%d["ms_per_step"], "rows/s %.3g"%d["value"], "pass frac %.3f"%d["roofline"]["frac"], d["kernel_ms_per_step"])
    except Exception as e: print(f, "failed", e)
PY
