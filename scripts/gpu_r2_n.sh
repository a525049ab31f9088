#!/usr/bin/env bash
set -uo pipefail
mkdir -p gpurun_out
( timeout 900 ./arrow_glue/test_glue 2>&1 | tail -6 )
for wl in sort_i64 filter; do ( timeout 600 ./arrow_glue/bench_glue $wl 100000000 5 ) > gpurun_out/glue_$wl.json 2> gpurun_out/glue_$wl.err; cat gpurun_out/glue_$wl.json; tail -2 gpurun_out/glue_$wl.err; done
( timeout 600 ./arrow_glue/bench_glue groupby_c3 200000000 3 ) > gpurun_out/glue_groupby_c3.json 2> gpurun_out/glue_groupby_c3.err; cat gpurun_out/glue_groupby_c3.json
T0=$(date +%s)
( timeout 1500 python -X faulthandler bench.py --steps 20 --warmup 5 ) > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
echo "default bench took $(( $(date +%s) - T0 )) s rc $?"; tail -3 gpurun_out/bench_default.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_default.json'))
print("main", d["config"]["workload"], "ms %.3f"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], "e2e ms %.2f"%d["e2e"]["ms_per_step"], "cpu", d["cpu_baseline"]["value"], d["cpu_baseline"]["sample"][:80])
for k,v in d.get("workloads",{}).items():
    if "error" in v: print(k, v); continue
    e=v.get("e2e") or {}; c=v.get("cpu_baseline") or {}
    print("%-12s ms %.3f rows/s %.3g frac %.3f | e2e ms %s | cpu %s"%(k, v["ms_per_step"], v["rows_per_s"], v["roofline"]["frac"] if v["roofline"] else -1, e.get("ms_per_step", e.get("error")), c.get("value")))
PY
