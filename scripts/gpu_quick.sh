#!/usr/bin/env bash
# a short GPU check: the test files (or -k expressions) given as arguments
set -uo pipefail
mkdir -p gpurun_out
( timeout 1200 python -m pytest "$@" -m gpu -x -q 2>&1 | tail -30 ) | tee gpurun_out/pytest_quick.log
