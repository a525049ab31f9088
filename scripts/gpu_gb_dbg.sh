#!/usr/bin/env bash
# phase-skipping timing experiment for the binned group-by kernel (results are wrong by construction)
for d in 0 8; do
  B200_GB_DBG=$d timeout 300 python bench.py --workload groupby_c3 --no-extras --no-cpu --no-e2e --steps 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('dbg=$d kernel ms %.3f'%d['kernel_ms_per_step']['groupby_consume'])"
done
