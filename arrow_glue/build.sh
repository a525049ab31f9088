#!/usr/bin/env bash
# Builds the Arrow-side glue against the UNMODIFIED reference where it lies (/root/reference):
#   arrow_glue/libarrow1_b200_arrow.a    RegisterCudaKernels / cuda::GroupBy / cuda::MakeGrouper
#   arrow_glue/test_glue                 GPU parity driver: CUDA registry vs the reference's CPU registry
# Links oracle/_ref/libarrow_ref.a (the reference's core + compute TUs, built by oracle/build_ref.sh),
# the reference's arrow_cuda TUs (gpu/cuda_{context,memory,internal}.cc, compiled here into
# arrow_glue/_build/) and libarrow1_b200.so.  No cmake, no reference build system.  On a box
# without /root/reference (the GPU box) the prebuilt files are used as shipped.
set -euo pipefail
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(dirname "$HERE")"
SRC="$REF/cpp/src"
OUT="$HERE/_build"
CUDA=${CUDA_HOME:-/usr/local/cuda}
if [ ! -d "$SRC/arrow" ]; then
  echo "arrow_glue/build.sh: $SRC not present (GPU box?) - using prebuilt files if any"; exit 0
fi
if [ ! -f "$ROOT/oracle/_ref/libarrow_ref.a" ]; then bash "$ROOT/oracle/build_ref.sh"; fi
mkdir -p "$OUT/include/arrow/gpu"
cat > "$OUT/include/arrow/gpu/cuda_version.h" <<'CFG'
#ifndef ARROW_GPU_CUDA_VERSION_H
#define ARROW_GPU_CUDA_VERSION_H
#define ARROW_CUDA_ABI_VERSION_MAJOR 12
#define ARROW_CUDA_ABI_VERSION_MINOR 9
#endif
CFG
CXXFLAGS="-std=c++11 -O2 -DNDEBUG -fPIC -w -include limits -include cstdint -DARROW_STATIC \
 -I$OUT/include -I$ROOT/oracle/_ref/include -I$SRC -I$CUDA/include -I$ROOT/include"
for f in cuda_context cuda_memory cuda_internal; do
  o="$OUT/gpu_$f.o"
  if [ ! -f "$o" ] || [ "$SRC/arrow/gpu/$f.cc" -nt "$o" ]; then
    g++ $CXXFLAGS -DARROW_EXPORTING -c "$SRC/arrow/gpu/$f.cc" -o "$o" &
  fi
done
wait
g++ $CXXFLAGS -c "$HERE/b200_arrow.cc" -o "$OUT/b200_arrow.o"
# the glue as a static archive (what a maintainer adds next to libarrow / libarrow_cuda); the parity
# driver links it with the reference objects statically, the C-ABI library and libcuda dynamically
rm -f "$HERE/libarrow1_b200_arrow.a"
ar rcs "$HERE/libarrow1_b200_arrow.a" "$OUT/b200_arrow.o"
g++ $CXXFLAGS -c "$HERE/test_glue.cc" -o "$OUT/test_glue.o"
g++ -o "$HERE/test_glue" "$OUT/test_glue.o" "$HERE/libarrow1_b200_arrow.a" "$OUT"/gpu_*.o \
  "$ROOT/oracle/_ref/libarrow_ref.a" -L"$ROOT/arrow1_b200" -larrow1_b200 -L"$CUDA/lib64/stubs" -lcuda \
  -Wl,-rpath,'$ORIGIN/../arrow1_b200' -lpthread -ldl
g++ $CXXFLAGS -c "$HERE/bench_glue.cc" -o "$OUT/bench_glue.o"
g++ -o "$HERE/bench_glue" "$OUT/bench_glue.o" "$HERE/libarrow1_b200_arrow.a" "$OUT"/gpu_*.o \
  "$ROOT/oracle/_ref/libarrow_ref.a" -L"$ROOT/arrow1_b200" -larrow1_b200 -L"$CUDA/lib64/stubs" -lcuda \
  -Wl,-rpath,'$ORIGIN/../arrow1_b200' -lpthread -ldl
echo "arrow_glue: $(ls -la "$HERE/libarrow1_b200_arrow.a" "$HERE/test_glue" | awk '{print $5, $9}' | tr '\n' ' ')"
