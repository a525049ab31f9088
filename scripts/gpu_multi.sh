#!/usr/bin/env bash
# multi-GPU pass (gpurun --gpus N): NCCL parity tests + the default bench line at N GPUs (with the §8e workloads)
set -uo pipefail
N=${1:-2}
mkdir -p gpurun_out
( timeout 900 python -m pytest tests/test_dist_gpu.py -m gpu -q 2>&1 | tail -15 ) > gpurun_out/pytest_dist_$N.log 2>&1
cat gpurun_out/pytest_dist_$N.log
( timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 10 --warmup 3 --no-cpu ) > gpurun_out/multi_${N}.json 2> gpurun_out/multi_${N}.err
tail -5 gpurun_out/multi_${N}.err
python - $N <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open('gpurun_out/multi_%s.json'%n))
    print("main", d["config"]["workload"], "ms %.3f"%d["ms_per_step"], "rows/s %.3g"%d["value"], "e2e", d["e2e"] and round(d["e2e"]["ms_per_step"],2), "check", d.get("self_check"))
    print(" kernels", {k:round(v,3) for k,v in d["kernel_ms_per_step"].items()})
    for k,v in d.get("workloads",{}).items():
        if v is None or "error" in v: print(k, v); continue
        print(k, "ms %.3f"%v["ms_per_step"], "rows/s %.3g"%v["rows_per_s"], v["self_check"], {a:round(b,3) for a,b in v["kernel_ms_per_step"].items()})
except Exception as e: print("failed", e)
PY
