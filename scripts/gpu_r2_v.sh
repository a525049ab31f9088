#!/usr/bin/env bash
# the no-alloc entry point: its test alone in a fresh process, then after the whole GPU suite
set -uo pipefail
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/pytest_noalloc_alone.log 2>&1; cat gpurun_out/pytest_noalloc_alone.log
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/pytest_gpu_all.log 2>&1; cat gpurun_out/pytest_gpu_all.log
