#!/usr/bin/env bash
# C4 launch list (after the same command ran clean) to see where the un-labelled time goes
set -uo pipefail
mkdir -p gpurun_out
CMD="python bench.py --workload groupby_c4 --no-extras --no-cpu --no-e2e --steps 2 --warmup 1"
( timeout 600 $CMD ) > gpurun_out/c4_plain.json 2> gpurun_out/c4_plain.err || { tail -5 gpurun_out/c4_plain.err; exit 1; }
python -c "
import json; d=json.load(open('gpurun_out/c4_plain.json')); print('c4 ms', d['ms_per_step'], {k:round(v,2) for k,v in d['kernel_ms_per_step'].items()})"
( timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"gb_|sort_|tail_|fill|memset" -c 600 --csv --log-file gpurun_out/launches_c4.csv $CMD ) > gpurun_out/ncu_c4.log 2>&1
python - <<'PY'
import csv, collections
rows = list(csv.reader(open('gpurun_out/launches_c4.csv')))
hdr = next(i for i, r in enumerate(rows) if r and r[0] == 'ID')
h = rows[hdr]
ki, vi = h.index('Kernel Name'), h.index('Metric Value')
seq = [(r[ki].split('(')[0][-60:], float(r[vi].replace(',', ''))) for r in rows[hdr+1:] if len(r) > vi]
# last step = the tail; print the last 70 launches in order
for k, v in seq[-45:]: print('%-60s %10.1f us' % (k, v / 1000 if v > 1e4 else v))
PY
