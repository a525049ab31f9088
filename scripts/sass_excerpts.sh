#!/usr/bin/env bash
# cuobjdump -sass of the dominant kernels -> profiles/r02_sass_<kernel>.txt (mnemonic census + the listing's first lines)
# and profiles/r02_sass_summary.txt.  Runs here (no GPU).
set -euo pipefail
cd "$(dirname "$0")/.."
LIB=arrow1_b200/libarrow1_b200.so
OUT=profiles/r02_sass_summary.txt
echo "# cuobjdump -sass $LIB (sm_100a), round 2: mnemonic census of the dominant kernels" > $OUT
echo "# whole library: $(cuobjdump -sass $LIB | grep -cE 'UBLKCP|UTMALDG|UTMASTG') bulk-copy / TMA instructions, $(cuobjdump -sass $LIB | grep -c 'LDG.E.128') LDG.E.128, $(cuobjdump -sass $LIB | grep -c 'LDG.E.64') LDG.E.64" >> $OUT
for k in sort_pass_kernelILb0ELb0ELb0 sort_pass_kernelILb1ELb0ELb0 sort_hist_kernelILb1 hw_runs_kernel hw_fix_kernel filter_apply_kernelImLb0ELb1 gb_bin32_kernel gb_hp_part_kernelILi1ELb0 gb_hp_agg_kernelILb1 gb_fused8_kernelILi4 gb_emit_kernel; do
  fn=$(cuobjdump -sass $LIB | grep -o "Function : [^ ]*$k[^ ]*" | head -1 | sed 's/Function : //')
  [ -z "$fn" ] && { echo "no $k" >> $OUT; continue; }
  f=profiles/r02_sass_$(echo $k | sed 's/IL.*//; s/Im.*//').txt
  [ "$k" = sort_pass_kernelILb1ELb0ELb0 ] && f=profiles/r02_sass_sort_pass_kernel_narrow.txt
  cuobjdump -sass -fun "$fn" $LIB 2>/dev/null | grep -E "^\s+/\*[0-9a-f]{4,}\*/" | sed -E 's#/\* 0x[0-9a-f]+ \*/##; s#\s+$##' > /tmp/sass_k.txt
  n=$(wc -l < /tmp/sass_k.txt)
  census=$(grep -oE '\b(LDG\.E[0-9A-Z.]*|STG\.E[0-9A-Z.]*|ATOMS\.[A-Z0-9.]*|LDS[0-9A-Z.]*|STS[0-9A-Z.]*|VOTE\.[A-Z.]*|REDUX[A-Z.]*|UBLKCP[A-Z.]*|SYNCS[A-Z.0-9]*|BAR\.SYNC[A-Z._]*|DADD|DMUL|ATOMG\.[A-Z0-9.]*|REDG\.[A-Z0-9.]*|LDL[0-9A-Z.]*|STL[0-9A-Z.]*)' /tmp/sass_k.txt | sort | uniq -c | sort -rn | awk '{printf "%s x%s, ", $2, $1}')
  echo "$k ($fn): $n instructions; $census" >> $OUT
  { echo "# $fn: $n SASS instructions; census: $census"; echo "# first 400 instructions:"; head -400 /tmp/sass_k.txt; } > $f
done
cat $OUT | cut -c1-300
