#!/usr/bin/env bash
# builds libarrow1_b200 variants with different sort-pass tile shapes into gpurun_variants/ (tuning sweep)
set -euo pipefail
cd "$(dirname "$0")/.."
mkdir -p variants
OBJ=arrow1_b200/csrc/_obj
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden -I include -I arrow1_b200/csrc"
for v in "$@"; do
  IFS=, read -r T I M <<< "$v"
  ( nvcc $FLAGS -DB200_SORT_THREADS=$T -DB200_SORT_ITEMS=$I -DB200_SORT_MINBLOCKS=$M -Xptxas -v -c arrow1_b200/csrc/sort.cu -o variants/sort_${T}_${I}_${M}.o 2> variants/sort_${T}_${I}_${M}.ptxas
    nvcc -shared -o variants/lib_${T}_${I}_${M}.so $OBJ/ctx.o $OBJ/filter.o $OBJ/groupby.o variants/sort_${T}_${I}_${M}.o -gencode arch=compute_100a,code=sm_100a
    echo "$v: $(grep -A2 'sort_pass_kernelILb0' variants/sort_${T}_${I}_${M}.ptxas | grep -oE '[0-9]+ bytes spill stores|Used [0-9]+ registers' | tr '\n' ' ')" ) &
done
wait
