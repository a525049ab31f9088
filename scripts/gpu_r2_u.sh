#!/usr/bin/env bash
set -uo pipefail
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_groupby_gpu.py tests/test_abi.py tests/test_q1_gpu.py tests/test_filter_gpu.py -m gpu -x -q 2>&1 | tail -25 )
( timeout 900 ./arrow_glue/test_glue 2>&1 ) > gpurun_out/test_glue.log 2>&1; grep -v "^ok" gpurun_out/test_glue.log | head -40
