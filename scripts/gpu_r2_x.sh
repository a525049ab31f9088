#!/usr/bin/env bash
set -uo pipefail
mkdir -p gpurun_out
rm -f gpurun_out/prof_r02_summary_*.txt
( timeout 1200 python -m pytest tests/test_sort_gpu.py tests/test_golden.py tests/test_abi.py tests/test_take_gpu.py tests/test_pipeline_gpu.py -m gpu -x -q 2>&1 | tail -6 )
for wl in sort_i64 sort_f64 take_sorted; do
  python bench.py --workload $wl --no-extras --no-cpu --no-e2e --steps 20 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$wl', 'ms %.3f'%d['ms_per_step'], {k:round(v,3) for k,v in d['kernel_ms_per_step'].items()}, d.get('gpu_launches'), d.get('plan'), 'frac %.3f'%d['roofline']['frac'])"
done
bash scripts/gpu_r2_prof.sh sort_i64 2>&1 | tail -5
