#!/usr/bin/env bash
# round-2 ncu captures: one --set full pass over one step of every hot-path workload (kernels named below);
# the summaries are made on the box, the report itself only travels back if it is small enough
set -uo pipefail
mkdir -p gpurun_out
WL=${1:-}
( timeout 600 python scripts/prof_all.py $WL ) > gpurun_out/prof_all_plain.log 2>&1 || { tail -20 gpurun_out/prof_all_plain.log; exit 1; }
cat gpurun_out/prof_all_plain.log
REP=/tmp/prof_r02_all
timeout 1700 ncu --set full --clock-control none --import-source on --profile-from-start off \
  -k regex:"sort_pass|sort_hist|hw_fix|tail_|filter_apply|gb_bin32|gb_hp_part|gb_hp_agg|gb_hp_hist|gb_fused8|gb_emit" \
  -o $REP python scripts/prof_all.py $WL > gpurun_out/prof_all_ncu.log 2>&1
echo "ncu rc $?"; tail -3 gpurun_out/prof_all_ncu.log; ls -la $REP.ncu-rep
ncu -i $REP.ncu-rep --page raw --csv > gpurun_out/prof_r02_all_raw.csv 2>/dev/null
# (the per-launch summaries are made afterwards, off the box: scripts/ncu_summary.py report.ncu-rep <launch>)
SZ=$(stat -c %s $REP.ncu-rep); if [ "$SZ" -lt 45000000 ]; then cp $REP.ncu-rep gpurun_out/prof_r02_all.ncu-rep; fi
