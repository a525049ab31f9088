#!/usr/bin/env bash
# multi-GPU pass (gpurun --gpus N): every workload's distributed form through bench.py under torchrun
set -uo pipefail
N=${1:-2}
mkdir -p gpurun_out
for wl in sort_i64 filter groupby_c3 groupby_c4; do
  rows=""
  if [ "$wl" = "groupby_c4" ]; then rows="--rows $((1000000000 / N))"; fi
  ( timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --workload $wl $rows --steps 5 --warmup 3 --no-extras --no-cpu ) > gpurun_out/multi_${N}_$wl.json 2> gpurun_out/multi_${N}_$wl.err
  python - "$wl" "$N" <<'PY'
import json,sys
wl,n=sys.argv[1],sys.argv[2]
try:
    d=json.loads([l for l in open('gpurun_out/multi_%s_%s.json'%(n,wl)) if l.startswith('{')][-1])
    print("%-12s gpus %d rows/s %.3g ms/step %.3f e2e rows/s %.3g"%(wl, d["n_gpus"], d["value"], d["ms_per_step"], (d.get("e2e") or {}).get("value") or 0))
except Exception as e:
    print(wl, "failed", e); print(open('gpurun_out/multi_%s_%s.err'%(n,wl)).read()[-1500:])
PY
done
