#!/usr/bin/env bash
# compare the two fast group-by variants (counting-sort tile vs direct CAS) on C3
set -uo pipefail
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_groupby_gpu.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -3 ) > gpurun_out/gb_pytest_sorted.log 2>&1
( B200_GB_DIRECT=1 timeout 600 python -m pytest tests/test_groupby_gpu.py tests/test_golden.py -m gpu -x -q 2>&1 | tail -3 ) > gpurun_out/gb_pytest_direct.log 2>&1
( timeout 300 python bench.py --workload groupby_c3 --no-extras --no-cpu --no-e2e --steps 5 ) > gpurun_out/gb_bench_sorted.json 2> gpurun_out/gb_bench.err
( B200_GB_DIRECT=1 timeout 300 python bench.py --workload groupby_c3 --no-extras --no-cpu --no-e2e --steps 5 ) > gpurun_out/gb_bench_direct.json 2>> gpurun_out/gb_bench.err
cat gpurun_out/gb_pytest_sorted.log gpurun_out/gb_pytest_direct.log
python - <<'PY'
import json
for v in ("sorted","direct"):
    try:
        d=json.load(open('gpurun_out/gb_bench_%s.json'%v))
        print(v, "ms/step %.2f"%d["ms_per_step"], "kernel ms", d["kernel_ms_per_step"].get("groupby_consume"), "frac %.3f"%d["roofline"]["frac"])
    except Exception as e: print(v, "failed", e)
PY
