#!/usr/bin/env bash
# round 2: group-by parity tests (incl. the partitioned high-cardinality path), C4 bench lines, launch list
set -uo pipefail
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_groupby_gpu.py tests/test_golden.py tests/test_repartition_gpu.py -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/gb_pytest.log 2>&1
cat gpurun_out/gb_pytest.log
for rows in 200000000 1000000000; do
  ( timeout 600 python bench.py --workload groupby_c4 --rows $rows --no-extras --no-cpu --no-e2e --steps 3 ) > gpurun_out/c4_$rows.json 2> gpurun_out/c4_$rows.err
  tail -3 gpurun_out/c4_$rows.err
  python - $rows <<'PY'
import json,sys
try:
    d=json.load(open('gpurun_out/c4_%s.json'%sys.argv[1]))
    print("rows", sys.argv[1], "ms/step %.2f"%d["ms_per_step"], "rows/s %.3g"%d["value"], "kernels", {k:round(v,2) for k,v in d["kernel_ms_per_step"].items()})
except Exception as e: print("failed", e)
PY
done
CMD="python bench.py --workload groupby_c4 --rows 200000000 --no-extras --no-cpu --no-e2e --steps 1 --warmup 3"
( timeout 300 $CMD > gpurun_out/c4_plain.log 2>&1 ) && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 120 --csv --log-file gpurun_out/c4_launches.csv $CMD > gpurun_out/c4_ncu.log 2>&1
echo "ncu rc $?"
python - <<'PY'
import csv,collections
try:
    rows=[r for r in csv.reader(open('gpurun_out/c4_launches.csv')) if len(r)>10 and r[0].isdigit()]
    agg=collections.OrderedDict()
    for r in rows:
        name=r[4].split('(')[0][-60:]; 
        try: v=float(r[-1].replace(',',''))
        except: continue
        a=agg.setdefault(name,[0,0.0]); a[0]+=1; a[1]+=v
    for k,(c,t) in agg.items(): print("%-62s n=%3d total %.3f ms"%(k,c,t/1e6))
except Exception as e: print("launch list failed", e)
PY
