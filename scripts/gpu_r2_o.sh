#!/usr/bin/env bash
set -uo pipefail
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_sort_gpu.py tests/test_golden.py tests/test_take_gpu.py tests/test_pipeline_gpu.py -m gpu -q 2>&1 | tail -5 )
for lib in arrow1_b200/libarrow1_b200.so variants/lib_notma.so; do
  for wl in sort_i64 sort_f64 sort_2keys; do
    B200_LIB=$lib python bench.py --workload $wl --no-extras --no-cpu --no-e2e --steps 20 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$lib', '$wl', 'ms %.3f'%d['ms_per_step'], {k:round(v,3) for k,v in d['kernel_ms_per_step'].items()})"
  done
done
( timeout 900 ./arrow_glue/test_glue 2>&1 | tail -4 )
for wl in sort_i64 filter; do ( timeout 600 ./arrow_glue/bench_glue $wl 100000000 5 ) > gpurun_out/glue_$wl.json 2> gpurun_out/glue_$wl.err; cat gpurun_out/glue_$wl.json; tail -2 gpurun_out/glue_$wl.err; done
( timeout 600 ./arrow_glue/bench_glue groupby_c3 200000000 3 ) > gpurun_out/glue_groupby_c3.json 2> gpurun_out/glue_groupby_c3.err; cat gpurun_out/glue_groupby_c3.json
python bench.py --workload groupby_c4 --rows 200000000 --no-extras --no-cpu --steps 3 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('c4 200M e2e ms', d['e2e']['ms_per_step'])"
