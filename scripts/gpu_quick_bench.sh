#!/usr/bin/env bash
# a short GPU comparison: bench.py workloads (arguments) under every library given in LIBS
set -uo pipefail
for lib in ${LIBS:-arrow1_b200/libarrow1_b200.so}; do
  for wl in "$@"; do
    B200_LIB=$lib python bench.py --workload $wl --no-extras --no-cpu --no-e2e --steps 20 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('$lib', '$wl', 'ms %.3f'%d['ms_per_step'], {k:round(v,3) for k,v in d['kernel_ms_per_step'].items()}, d.get('plan'))"
  done
done
