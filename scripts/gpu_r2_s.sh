#!/usr/bin/env bash
set -uo pipefail
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_groupby_gpu.py tests/test_full_size_gpu.py tests/test_repartition_gpu.py -m gpu -x -q -k "groupby or config4 or c4 or C4 or repartition" 2>&1 | tail -5 )
for lib in variants/lib_agg_0.so variants/lib_agg_a.so arrow1_b200/libarrow1_b200.so; do
  B200_LIB=$lib python bench.py --workload groupby_c4 --no-extras --no-cpu --no-e2e --steps 3 --warmup 2 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('c4 [$lib] ms %.2f'%d['ms_per_step'], {k:round(v,2) for k,v in d['kernel_ms_per_step'].items()})"
done
