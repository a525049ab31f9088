#!/usr/bin/env bash
# targeted capture: gb_bin32_kernel only (C3 shape, 200 M rows), report brought back for source-level reading
set -uo pipefail
mkdir -p gpurun_out
( timeout 300 python scripts/prof_all.py groupby_c3 ) > gpurun_out/prof_c3_plain.log 2>&1 || { tail -5 gpurun_out/prof_c3_plain.log; exit 1; }
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:gb_bin32 -c 1 \
  -o gpurun_out/prof_c3 python scripts/prof_all.py groupby_c3 > gpurun_out/prof_c3_ncu.log 2>&1
echo "ncu rc $?"; ls -la gpurun_out/prof_c3.ncu-rep
