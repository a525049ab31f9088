#!/usr/bin/env bash
set -uo pipefail
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_sort_gpu.py tests/test_golden.py tests/test_take_gpu.py -m gpu -q 2>&1 | tail -5 )
for lib in "" variants/lib_w3_n3.so variants/lib_w2_n3.so variants/lib_w2_n2.so; do
  for wl in sort_i64 sort_f64; do
    B200_LIB=${lib:-arrow1_b200/libarrow1_b200.so} python bench.py --workload $wl --no-extras --no-cpu --no-e2e --steps 20 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('${lib:-default}', '$wl', 'ms %.3f'%d['ms_per_step'], {k:round(v,3) for k,v in d['kernel_ms_per_step'].items()})"
  done
done
timeout 900 python -X faulthandler bench.py --workload groupby_c4 --no-extras --steps 3 > gpurun_out/c4_full.json 2> gpurun_out/c4_full.err; echo rc $?; tail -3 gpurun_out/c4_full.err; python -c "
import json; d=json.load(open('gpurun_out/c4_full.json')); print('c4', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], 'cpu', d['cpu_baseline'])"
