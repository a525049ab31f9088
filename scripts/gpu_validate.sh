#!/usr/bin/env bash
# round-2 validation pass: whole GPU suite, cub yardstick, default bench + reference arm, launch list under ncu
set -uo pipefail
mkdir -p gpurun_out
T0=$(date +%s)
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 ) > gpurun_out/pytest_gpu.log 2>&1
cat gpurun_out/pytest_gpu.log; echo "gpu suite took $(( $(date +%s) - T0 )) s"
( timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3 )
( timeout 300 ./scripts/ubench/cub_sort 100000000 10 ) > gpurun_out/cub_sort.json 2>&1; cat gpurun_out/cub_sort.json
T0=$(date +%s)
( timeout 900 python bench.py --impl reference --steps 3 --warmup 1 ) > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
echo "reference arm took $(( $(date +%s) - T0 )) s"; cut -c1-400 gpurun_out/bench_ref.json
T0=$(date +%s)
( timeout 1500 python bench.py ) > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err
echo "default bench took $(( $(date +%s) - T0 )) s"; tail -3 gpurun_out/bench_default.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_default.json'))
print("main", d["config"]["workload"], "ms %.3f"%d["ms_per_step"], "frac %.3f"%d["roofline"]["frac"], "e2e ms %.2f"%d["e2e"]["ms_per_step"], "cpu", d["cpu_baseline"]["value"], "launches", d.get("gpu_launches"), d.get("clocks"))
for k,v in d.get("workloads",{}).items():
    if "error" in v: print(k, v); continue
    e=v.get("e2e") or {}; c=v.get("cpu_baseline") or {}
    print("%-12s ms %.3f rows/s %.3g frac %.3f | e2e ms %s | cpu %s"%(k, v["ms_per_step"], v["rows_per_s"], v["roofline"]["frac"] if v["roofline"] else -1, e.get("ms_per_step", e.get("error")), c.get("value")))
PY
# launch list of the default bench command's device part (after the same command exited 0 above)
( timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_sort.csv \
    python bench.py --steps 2 --warmup 1 --no-extras --no-cpu --no-e2e ) > gpurun_out/ncu_launches.log 2>&1
tail -2 gpurun_out/ncu_launches.log; wc -l gpurun_out/launches_sort.csv
