#!/usr/bin/env bash
# what the driver runs at round end, minus the bench: the GPU suite and smoke()
set -uo pipefail
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 ) | tee gpurun_out/pytest_gpu.log
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 )
