#!/usr/bin/env bash
set -uo pipefail
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_sort_gpu.py tests/test_golden.py tests/test_full_size_gpu.py tests/test_take_gpu.py tests/test_groupby_gpu.py -m gpu -x -q -k "sort or golden or config2 or take or first_rows or groupings" 2>&1 | tail -4 )
for wl in sort_i64 sort_f64 sort_2keys groupby_c4; do
  python bench.py --workload $wl --no-extras --no-cpu --no-e2e --steps 10 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$wl', 'ms %.3f'%d['ms_per_step'], {k:round(v,3) for k,v in d['kernel_ms_per_step'].items()}, d.get('gpu_launches'))"
done
