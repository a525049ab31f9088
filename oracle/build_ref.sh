#!/usr/bin/env bash
# Builds the UNMODIFIED reference (Arrow C++ 5.0.0-SNAPSHOT under /root/reference) core+compute
# translation units straight from where they lie, with plain g++ (no cmake, no reference build
# system), into oracle/_ref/libarrow_ref.a.  Test infrastructure only: nothing on the product
# path links this.  The TU list follows cpp/src/arrow/CMakeLists.txt:122-215 (ARROW_SRCS) and
# :356-395 (ARROW_COMPUTE) minus hdfs/compression/uri which the hot path never reaches.
# The only generated header, arrow/util/config.h (util/config.h.cmake), is a dozen version
# #defines and is written by hand below.
set -euo pipefail
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
SRC="$REF/cpp/src"
JOBS=${JOBS:-$(nproc)}
if [ ! -d "$SRC/arrow" ]; then
  echo "build_ref: $SRC not present (GPU box?) - using prebuilt $OUT if any"; exit 0
fi
mkdir -p "$OUT/obj" "$OUT/include/arrow/util"
cat > "$OUT/include/arrow/util/config.h" <<'CFG'
#define ARROW_VERSION_MAJOR 5
#define ARROW_VERSION_MINOR 0
#define ARROW_VERSION_PATCH 0
#define ARROW_VERSION ((ARROW_VERSION_MAJOR * 1000) + ARROW_VERSION_MINOR) * 1000 + ARROW_VERSION_PATCH
#define ARROW_VERSION_STRING "5.0.0-SNAPSHOT"
#define ARROW_SO_VERSION "500"
#define ARROW_FULL_SO_VERSION "500.0.0"
#define ARROW_CXX_COMPILER_ID "GNU"
#define ARROW_CXX_COMPILER_VERSION "13"
#define ARROW_CXX_COMPILER_FLAGS ""
#define ARROW_GIT_ID ""
#define ARROW_GIT_DESCRIPTION ""
#define ARROW_PACKAGE_KIND ""
#define ARROW_COMPUTE
#define ARROW_USE_NATIVE_INT128
CFG
SRCS="
array/array_base.cc array/array_binary.cc array/array_decimal.cc array/array_dict.cc
array/array_nested.cc array/array_primitive.cc array/builder_adaptive.cc array/builder_base.cc
array/builder_binary.cc array/builder_decimal.cc array/builder_dict.cc array/builder_nested.cc
array/builder_primitive.cc array/builder_union.cc array/concatenate.cc array/data.cc array/diff.cc
array/util.cc array/validate.cc builder.cc buffer.cc chunked_array.cc compare.cc config.cc datum.cc
device.cc extension_type.cc memory_pool.cc pretty_print.cc record_batch.cc result.cc scalar.cc
sparse_tensor.cc status.cc table.cc table_builder.cc tensor.cc tensor/coo_converter.cc
tensor/csf_converter.cc tensor/csx_converter.cc type.cc visitor.cc
io/interfaces.cc io/memory.cc c/bridge.cc
util/basic_decimal.cc util/bit_block_counter.cc util/bit_run_reader.cc util/bit_util.cc
util/bitmap.cc util/bitmap_builders.cc util/bitmap_ops.cc util/bpacking.cc util/cancel.cc
util/cpu_info.cc util/decimal.cc util/delimiting.cc util/formatting.cc util/future.cc
util/int_util.cc util/io_util.cc util/logging.cc util/key_value_metadata.cc util/memory.cc
util/mutex.cc util/string.cc util/string_builder.cc util/task_group.cc util/tdigest.cc
util/thread_pool.cc util/time.cc util/trie.cc util/utf8.cc util/value_parsing.cc
vendored/base64.cpp vendored/datetime/tz.cpp
vendored/double-conversion/bignum.cc vendored/double-conversion/double-conversion.cc
vendored/double-conversion/bignum-dtoa.cc vendored/double-conversion/fast-dtoa.cc
vendored/double-conversion/cached-powers.cc vendored/double-conversion/fixed-dtoa.cc
vendored/double-conversion/diy-fp.cc vendored/double-conversion/strtod.cc
compute/api_aggregate.cc compute/api_scalar.cc compute/api_vector.cc compute/cast.cc compute/exec.cc
compute/exec/exec_plan.cc compute/exec/expression.cc compute/function.cc compute/kernel.cc
compute/registry.cc compute/kernels/aggregate_basic.cc compute/kernels/aggregate_mode.cc
compute/kernels/aggregate_quantile.cc compute/kernels/aggregate_tdigest.cc
compute/kernels/aggregate_var_std.cc compute/kernels/codegen_internal.cc
compute/kernels/hash_aggregate.cc compute/kernels/scalar_arithmetic.cc
compute/kernels/scalar_boolean.cc compute/kernels/scalar_cast_boolean.cc
compute/kernels/scalar_cast_internal.cc compute/kernels/scalar_cast_nested.cc
compute/kernels/scalar_cast_numeric.cc compute/kernels/scalar_cast_string.cc
compute/kernels/scalar_cast_temporal.cc compute/kernels/scalar_compare.cc
compute/kernels/scalar_nested.cc compute/kernels/scalar_set_lookup.cc
compute/kernels/scalar_string.cc compute/kernels/scalar_validity.cc
compute/kernels/scalar_fill_null.cc compute/kernels/util_internal.cc compute/kernels/vector_hash.cc
compute/kernels/vector_nested.cc compute/kernels/vector_selection.cc compute/kernels/vector_sort.cc
"
CXXFLAGS="-std=c++11 -O3 -DNDEBUG -fPIC -w -mpopcnt -msse4.2 -include limits -include cstdint \
 -DARROW_STATIC -DARROW_EXPORTING -DURI_STATIC_BUILD \
 -I$OUT/include -I$SRC -I$SRC/arrow/vendored"
compile_one() {
  f="$1"; o="$OUT/obj/$(echo "$f" | tr '/' '_').o"
  if [ ! -f "$o" ] || [ "$SRC/arrow/$f" -nt "$o" ]; then
    g++ $CXXFLAGS -c "$SRC/arrow/$f" -o "$o" || { echo "FAILED $f"; return 1; }
  fi
}
export -f compile_one; export OUT SRC CXXFLAGS
echo $SRCS | tr ' ' '\n' | xargs -P "$JOBS" -I{} bash -c 'compile_one {}'
rm -f "$OUT/libarrow_ref.a"
ar rcs "$OUT/libarrow_ref.a" "$OUT"/obj/*.o
echo "build_ref: $(ls -la "$OUT/libarrow_ref.a")"
