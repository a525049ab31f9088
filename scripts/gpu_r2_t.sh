#!/usr/bin/env bash
set -uo pipefail
mkdir -p gpurun_out
( timeout 1200 python -m pytest tests/test_sort_gpu.py tests/test_golden.py tests/test_take_gpu.py tests/test_full_size_gpu.py -m gpu -x -q -k "sort or golden or take or config2" 2>&1 | tail -5 )
for lib in variants/lib_match0.so arrow1_b200/libarrow1_b200.so; do
  for wl in sort_i64 sort_f64 sort_2keys; do
    B200_LIB=$lib python bench.py --workload $wl --no-extras --no-cpu --no-e2e --steps 20 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$lib', '$wl', 'ms %.3f'%d['ms_per_step'], {k:round(v,3) for k,v in d['kernel_ms_per_step'].items()})"
  done
done
( B200_LIB=variants/lib_hp_i4.so timeout 900 python -m pytest tests/test_groupby_gpu.py -m gpu -x -q -k "partitioned" 2>&1 | tail -3 )
for lib in arrow1_b200/libarrow1_b200.so variants/lib_hp_i4.so; do
  B200_LIB=$lib python bench.py --workload groupby_c4 --no-extras --no-cpu --no-e2e --steps 3 --warmup 2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('c4 [$lib] ms %.2f'%d['ms_per_step'], {k:round(v,2) for k,v in d['kernel_ms_per_step'].items()})"
done
