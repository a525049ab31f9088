#!/usr/bin/env bash
# builds alternative libraries for a tuning sweep: scripts/build_variants.sh name "-DX=1 -DY=2" [source.cu]
# -> arrow1_b200/_variants/lib_<name>.so (git-ignored like every .so; travels with gpurun)
set -euo pipefail
NAME=$1; DEFS=$2; SRC=${3:-groupby.cu}
cd "$(dirname "$0")/.."
mkdir -p arrow1_b200/_variants /tmp/variants_obj
python -c "from arrow1_b200 import build; build.build()" >/dev/null
OBJ=/tmp/variants_obj/${NAME}_${SRC%.cu}.o
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-fvisibility=hidden \
  -I include -I arrow1_b200/csrc $DEFS -c arrow1_b200/csrc/$SRC -o $OBJ
OTHERS=$(ls arrow1_b200/csrc/_obj/*.o 2>/dev/null | grep -v "/${SRC%.cu}.o" || true)
nvcc -shared -o arrow1_b200/_variants/lib_$NAME.so $OBJ $OTHERS -gencode arch=compute_100a,code=sm_100a
ls -la arrow1_b200/_variants/lib_$NAME.so
