#!/usr/bin/env bash
set -uo pipefail
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_sort_gpu.py tests/test_golden.py tests/test_take_gpu.py tests/test_abi.py tests/test_pipeline_gpu.py -m gpu -x -q 2>&1 | tail -25 )
( timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2 )
python bench.py --workload sort_2keys --no-extras --no-cpu --no-e2e --steps 10 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('sort_2keys ms %.3f frac %.3f'%(d['ms_per_step'], d['roofline']['frac']), d['roofline']['launches_per_step'])"
