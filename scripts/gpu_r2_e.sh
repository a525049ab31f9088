#!/usr/bin/env bash
set -uo pipefail
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests/test_groupby_gpu.py tests/test_q1_gpu.py tests/test_filter_gpu.py tests/test_repartition_gpu.py -m gpu -q 2>&1 | tail -30 ) > gpurun_out/pytest_gpu.log 2>&1
cat gpurun_out/pytest_gpu.log
for wl in q1_fused groupby_c4; do
  ( timeout 600 python bench.py --workload $wl --no-extras --no-cpu --no-e2e --steps 5 ) > gpurun_out/b_$wl.json 2> gpurun_out/b_$wl.err
  tail -2 gpurun_out/b_$wl.err
  python - $wl <<'PY'
import json,sys
try:
    d=json.load(open('gpurun_out/b_%s.json'%sys.argv[1]))
    print(sys.argv[1], "ms/step %.3f"%d["ms_per_step"], "rows/s %.3g"%d["value"], "frac", d["roofline"] and round(d["roofline"]["frac"],3), "kernels", {k:round(v,3) for k,v in d["kernel_ms_per_step"].items()})
except Exception as e: print(sys.argv[1], "failed", e)
PY
done
CMD="python bench.py --workload groupby_c4 --rows 200000000 --no-extras --no-cpu --no-e2e --steps 1 --warmup 3"
( timeout 300 $CMD > gpurun_out/c4_plain.log 2>&1 ) && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 150 --csv --log-file gpurun_out/c4_launches.csv $CMD > gpurun_out/c4_ncu.log 2>&1
python - <<'PY'
import csv,collections
try:
    rows=[r for r in csv.reader(open('gpurun_out/c4_launches.csv')) if len(r)>10 and r[0].isdigit()]
    agg=collections.OrderedDict()
    for r in rows:
        name=r[4].split('(')[0][-50:]
        try: v=float(r[-1].replace(',',''))
        except: continue
        a=agg.setdefault(name,[0,0.0]); a[0]+=1; a[1]+=v
    for k,(c,t) in agg.items(): print("%-52s n=%3d total %.3f ms"%(k,c,t/1e6))
except Exception as e: print("launch list failed", e)
PY
