// TEST INFRASTRUCTURE ONLY — never linked or loaded by the product path.
//
// Thin extern "C" driver over the UNMODIFIED reference (Arrow C++ 5.0.0-SNAPSHOT compiled from
// /root/reference by oracle/build_ref.sh into oracle/_ref/libarrow_ref.a).  It wraps caller-owned
// HOST buffers as Arrow arrays (zero copy), calls the reference's own public entry points
//   compute::Filter            (compute/api_vector.cc:94-99)
//   compute::SortIndices       (compute/api_vector.cc:48-53)
//   compute::internal::GroupBy (compute/kernels/hash_aggregate.cc:890-966)
//   compute::internal::Grouper (compute/kernels/hash_aggregate.cc:313-437)
// and copies the results into caller-owned buffers.  Column descriptors reuse the b200_col
// layout of include/arrow1_b200.h with host pointers.  Used to (1) pin oracle/oracle.cc,
// (2) act as parity checker in tests/, (3) be timed as bench.py's cpu_baseline ("reference").
#include <algorithm>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "arrow/api.h"
#include "arrow/compute/api.h"
#include "arrow/compute/exec.h"
#include "arrow/util/bit_util.h"
#include "arrow/util/bitmap_ops.h"
#include "arrow1_b200.h"

namespace {

using namespace arrow;  // NOLINT

thread_local std::string g_error;

int Fail(const Status& st) {
  g_error = st.ToString();
  return static_cast<int>(st.code());
}

std::shared_ptr<DataType> TypeOf(int t) {
  switch (t) {
    case B200_BOOL: return boolean();
    case B200_UINT8: return uint8();
    case B200_INT8: return int8();
    case B200_UINT16: return uint16();
    case B200_INT16: return int16();
    case B200_UINT32: return uint32();
    case B200_INT32: return int32();
    case B200_UINT64: return uint64();
    case B200_INT64: return int64();
    case B200_FLOAT: return float32();
    case B200_DOUBLE: return float64();
    case B200_BINARY: return binary();
  }
  return nullptr;
}

int BitsOf(int t) {
  switch (t) {
    case B200_BOOL: return 1;
    case B200_UINT8: case B200_INT8: return 8;
    case B200_UINT16: case B200_INT16: return 16;
    case B200_UINT32: case B200_INT32: case B200_FLOAT: return 32;
    default: return 64;
  }
}

std::shared_ptr<ArrayData> Wrap(const b200_col& c) {
  const int bits = BitsOf(c.type);
  const int64_t n = c.offset + c.length;
  std::shared_ptr<Buffer> validity;
  if (c.validity) validity = std::make_shared<Buffer>(static_cast<const uint8_t*>(c.validity), BitUtil::BytesForBits(n));
  if (c.type == B200_BINARY) {   // data = int32 offsets, extra = value bytes
    const int32_t* offs = static_cast<const int32_t*>(c.data);
    auto offsets = std::make_shared<Buffer>(static_cast<const uint8_t*>(c.data), (n + 1) * 4);
    static const uint8_t kNoBytes[1] = {0};
    auto bytes = std::make_shared<Buffer>(c.extra ? static_cast<const uint8_t*>(c.extra) : kNoBytes, offs[n]);
    return ArrayData::Make(binary(), c.length, {validity, offsets, bytes}, c.validity ? c.null_count : 0, c.offset);
  }
  auto data = std::make_shared<Buffer>(static_cast<const uint8_t*>(c.data),
                                       bits == 1 ? BitUtil::BytesForBits(n) : n * (bits / 8));
  return ArrayData::Make(TypeOf(c.type), c.length, {validity, data}, c.validity ? c.null_count : 0, c.offset);
}

// copies an Arrow primitive array into caller buffers (offset 0 on the output side)
void CopyOut(const ArrayData& a, b200_out_col* out) {
  out->length = a.length;
  out->null_count = a.null_count.load();
  if (a.type->id() == Type::BINARY || a.type->id() == Type::STRING) {
    // offsets rebased to 0 into out->data (length + 1 entries), the bytes into out->extra
    const int32_t* offs = a.GetValues<int32_t>(1);
    int32_t* dst = static_cast<int32_t*>(out->data);
    for (int64_t i = 0; i <= a.length; ++i) dst[i] = offs[i] - offs[0];
    if (out->extra && a.buffers[2] && dst[a.length] > 0) std::memcpy(out->extra, a.buffers[2]->data() + offs[0], size_t(dst[a.length]));
    if (out->validity) {
      std::memset(out->validity, 0, BitUtil::BytesForBits(a.length));
      if (a.buffers[0]) arrow::internal::CopyBitmap(a.buffers[0]->data(), a.offset, a.length, static_cast<uint8_t*>(out->validity), 0);
      else BitUtil::SetBitsTo(static_cast<uint8_t*>(out->validity), 0, a.length, true);
    }
    return;
  }
  const int bits = arrow::internal::checked_cast<const FixedWidthType&>(*a.type).bit_width();
  if (a.length == 0) return;
  if (out->data) {
    if (bits == 1) {
      std::memset(out->data, 0, BitUtil::BytesForBits(a.length));
      arrow::internal::CopyBitmap(a.buffers[1]->data(), a.offset, a.length, static_cast<uint8_t*>(out->data), 0);
    } else {
      std::memcpy(out->data, a.buffers[1]->data() + a.offset * (bits / 8), a.length * (bits / 8));
    }
  }
  if (out->validity) {
    if (a.buffers[0]) {
      std::memset(out->validity, 0, BitUtil::BytesForBits(a.length));
      arrow::internal::CopyBitmap(a.buffers[0]->data(), a.offset, a.length, static_cast<uint8_t*>(out->validity), 0);
    } else {
      std::memset(out->validity, 0, BitUtil::BytesForBits(a.length));
      BitUtil::SetBitsTo(static_cast<uint8_t*>(out->validity), 0, a.length, true);
    }
  }
}

}  // namespace

extern "C" {

B200_API const char* ref_last_error() { return g_error.c_str(); }

B200_API const char* ref_version() { return "arrow-5.0.0-SNAPSHOT (reference, g++ -O3 -mpopcnt -msse4.2)"; }

// compute::Filter on (values, mask).  out buffers sized for values->length elements.
// *has_validity reports whether the reference allocated a validity bitmap.
B200_API int ref_filter(const b200_col* values, const b200_col* mask, int null_selection, b200_out_col* out,
                        int* has_validity) {
  compute::FilterOptions opts(null_selection ? compute::FilterOptions::EMIT_NULL : compute::FilterOptions::DROP);
  compute::ExecContext ctx;
  auto res = compute::Filter(Datum(Wrap(*values)), Datum(Wrap(*mask)), opts, &ctx);
  if (!res.ok()) return Fail(res.status());
  auto arr = res.ValueOrDie().array();
  if (has_validity) *has_validity = arr->buffers[0] != nullptr;
  CopyOut(*arr, out);
  return 0;
}

// only runs the kernel, keeps the output inside Arrow (for timing without the copy-out)
B200_API int ref_filter_timed(const b200_col* values, const b200_col* mask, int null_selection, int64_t* out_len) {
  compute::FilterOptions opts(null_selection ? compute::FilterOptions::EMIT_NULL : compute::FilterOptions::DROP);
  compute::ExecContext ctx;
  auto res = compute::Filter(Datum(Wrap(*values)), Datum(Wrap(*mask)), opts, &ctx);
  if (!res.ok()) return Fail(res.status());
  if (out_len) *out_len = res.ValueOrDie().length();
  return 0;
}

// compute::Take(values, indices, TakeOptions(boundscheck)).  out buffers sized for indices->length elements; the
// reference always allocates the validity bitmap (PrimitiveTake, vector_selection.cc:493-506).
B200_API int ref_take(const b200_col* values, const b200_col* indices, int boundscheck, b200_out_col* out) {
  compute::TakeOptions opts(boundscheck != 0);
  compute::ExecContext ctx;
  auto res = compute::Take(Datum(Wrap(*values)), Datum(Wrap(*indices)), opts, &ctx);
  if (!res.ok()) return Fail(res.status());
  CopyOut(*res.ValueOrDie().array(), out);
  return 0;
}

// compute::SortIndices(Array, order); index_base is added afterwards (ChunkedArraySorter offset)
B200_API int ref_sort_indices(const b200_col* values, int order, uint64_t index_base, uint64_t* out) {
  compute::ExecContext ctx;
  auto arr = MakeArray(Wrap(*values));
  auto res = compute::SortIndices(*arr, order ? compute::SortOrder::Descending : compute::SortOrder::Ascending, &ctx);
  if (!res.ok()) return Fail(res.status());
  auto idx = res.ValueOrDie()->data();
  const uint64_t* p = idx->GetValues<uint64_t>(1);
  if (out) {
    for (int64_t i = 0; i < idx->length; ++i) out[i] = p[i] + index_base;
  }
  return 0;
}

// compute::SortIndices(Datum(RecordBatch | Table), SortOptions{sort_keys}): keys[k] becomes column "k<k>".
// as_table != 0 wraps the columns in a Table of `n_chunks` chunks (MultipleKeyTableSorter), else a RecordBatch
// (RadixRecordBatchSorter up to 8 keys, MultipleKeyRecordBatchSorter beyond; vector_sort.cc:1632-1711).
B200_API int ref_sort_indices_multi(const b200_col* keys, const int* orders, int n_keys, int as_table, int n_chunks,
                                    uint64_t* out) {
  compute::ExecContext ctx;
  std::vector<std::shared_ptr<Field>> fields;
  std::vector<std::shared_ptr<Array>> columns;
  std::vector<compute::SortKey> sort_keys;
  for (int k = 0; k < n_keys; ++k) {
    const std::string name = "k" + std::to_string(k);
    columns.push_back(MakeArray(Wrap(keys[k])));
    fields.push_back(field(name, columns.back()->type()));
    sort_keys.emplace_back(name, orders[k] ? compute::SortOrder::Descending : compute::SortOrder::Ascending);
  }
  compute::SortOptions options(sort_keys);
  const int64_t n = n_keys > 0 ? keys[0].length : 0;
  Datum input;
  if (as_table) {
    std::vector<std::shared_ptr<ChunkedArray>> chunked;
    const int64_t step = std::max<int64_t>(1, (n + std::max(n_chunks, 1) - 1) / std::max(n_chunks, 1));
    for (const auto& col : columns) {
      ArrayVector pieces;
      for (int64_t at = 0; at < n; at += step) pieces.push_back(col->Slice(at, std::min(step, n - at)));
      chunked.push_back(std::make_shared<ChunkedArray>(pieces, col->type()));
    }
    input = Datum(Table::Make(schema(fields), chunked, n));
  } else {
    input = Datum(RecordBatch::Make(schema(fields), n, columns));
  }
  auto res = compute::SortIndices(input, options, &ctx);
  if (!res.ok()) return Fail(res.status());
  auto idx = res.ValueOrDie()->data();
  const uint64_t* p = idx->GetValues<uint64_t>(1);
  if (out) std::memcpy(out, p, size_t(idx->length) * 8);
  return 0;
}

// compute::internal::GroupBy, single batch (the behaviour the reference's tests pin).
// hash_mean does not exist in the reference: it is produced as double(sum)/count from the
// reference's own hash_sum and hash_count columns (MeanImpl::Finalize, aggregate_basic_internal.h:99-109).
// Output layout == b200_groupby_finalize.  out_ids (nullable): Grouper::Consume ids per row.
B200_API int ref_groupby(int n_keys, const b200_col* keys, int n_aggs, const b200_agg_spec* aggs,
                         const b200_col* args, b200_out_col* out_aggs, b200_out_col* out_keys,
                         uint32_t* out_num_groups, uint32_t* out_ids) {
  compute::ExecContext ctx;
  std::vector<Datum> key_datums, arg_datums;
  for (int k = 0; k < n_keys; ++k) key_datums.emplace_back(Wrap(keys[k]));
  std::vector<compute::internal::Aggregate> specs;
  std::vector<std::unique_ptr<compute::FunctionOptions>> keep;
  // plan: each requested agg maps to one or two reference aggregates
  struct Map { int first; int second; };
  std::vector<Map> maps;
  for (int j = 0; j < n_aggs; ++j) {
    Map m{static_cast<int>(specs.size()), -1};
    switch (aggs[j].kind) {
      case B200_HASH_COUNT: {
        keep.emplace_back(new compute::CountOptions(aggs[j].option ? compute::CountOptions::COUNT_NULL
                                                                    : compute::CountOptions::COUNT_NON_NULL));
        specs.push_back({"hash_count", keep.back().get()});
        arg_datums.emplace_back(Wrap(args[j]));
        break;
      }
      case B200_HASH_SUM:
        specs.push_back({"hash_sum", nullptr});
        arg_datums.emplace_back(Wrap(args[j]));
        break;
      case B200_HASH_MIN_MAX: {
        keep.emplace_back(new compute::MinMaxOptions(aggs[j].option ? compute::MinMaxOptions::EMIT_NULL
                                                                     : compute::MinMaxOptions::SKIP));
        specs.push_back({"hash_min_max", keep.back().get()});
        arg_datums.emplace_back(Wrap(args[j]));
        break;
      }
      case B200_HASH_MEAN: {
        specs.push_back({"hash_sum", nullptr});
        arg_datums.emplace_back(Wrap(args[j]));
        keep.emplace_back(new compute::CountOptions(compute::CountOptions::COUNT_NON_NULL));
        m.second = static_cast<int>(specs.size());
        specs.push_back({"hash_count", keep.back().get()});
        arg_datums.emplace_back(Wrap(args[j]));
        break;
      }
      default:
        g_error = "bad agg kind";
        return B200_INVALID;
    }
    maps.push_back(m);
  }
  uint32_t num_groups = 0;
  if (n_aggs > 0) {
    auto res = compute::internal::GroupBy(arg_datums, key_datums, specs, &ctx);
    if (!res.ok()) return Fail(res.status());
    auto st = res.ValueOrDie().array();
    num_groups = static_cast<uint32_t>(st->length);
    int o = 0;
    for (int j = 0; j < n_aggs; ++j) {
      const auto& col = *st->child_data[maps[j].first];
      switch (aggs[j].kind) {
        case B200_HASH_COUNT:
        case B200_HASH_SUM:
          CopyOut(col, &out_aggs[o++]);
          break;
        case B200_HASH_MIN_MAX: {
          // struct<min,max>: children share the validity bitmap (hash_aggregate.cc:741-757)
          CopyOut(*col.child_data[0], &out_aggs[o++]);
          CopyOut(*col.child_data[1], &out_aggs[o++]);
          break;
        }
        case B200_HASH_MEAN: {
          const auto& cnt = *st->child_data[maps[j].second];
          const int64_t* counts = cnt.GetValues<int64_t>(1);
          b200_out_col* out = &out_aggs[o++];
          out->length = num_groups;
          double* dst = static_cast<double*>(out->data);
          uint8_t* val = static_cast<uint8_t*>(out->validity);
          if (val) std::memset(val, 0, BitUtil::BytesForBits(num_groups));
          int64_t nulls = 0;
          for (uint32_t g = 0; g < num_groups; ++g) {
            if (counts[g] == 0) {
              ++nulls;
              dst[g] = 0;
              continue;
            }
            double s;
            switch (col.type->id()) {
              case Type::INT64: s = static_cast<double>(col.GetValues<int64_t>(1)[g]); break;
              case Type::UINT64: s = static_cast<double>(col.GetValues<uint64_t>(1)[g]); break;
              default: s = col.GetValues<double>(1)[g]; break;
            }
            dst[g] = s / static_cast<double>(counts[g]);
            if (val) BitUtil::SetBit(val, g);
          }
          out->null_count = nulls;
          break;
        }
      }
    }
    for (int k = 0; k < n_keys; ++k) {
      const auto& col = *st->child_data[specs.size() + k];
      CopyOut(col, &out_keys[k]);
    }
  }
  if (out_ids || n_aggs == 0) {
    std::vector<ValueDescr> descrs;
    for (int k = 0; k < n_keys; ++k) descrs.emplace_back(TypeOf(keys[k].type), ValueDescr::ARRAY);
    auto gr = compute::internal::Grouper::Make(descrs, &ctx);
    if (!gr.ok()) return Fail(gr.status());
    auto grouper = std::move(gr).ValueOrDie();
    auto batch = compute::ExecBatch::Make(key_datums);
    if (!batch.ok()) return Fail(batch.status());
    auto ids = grouper->Consume(batch.ValueOrDie());
    if (!ids.ok()) return Fail(ids.status());
    auto id_arr = ids.ValueOrDie().array();
    if (out_ids) std::memcpy(out_ids, id_arr->GetValues<uint32_t>(1), id_arr->length * sizeof(uint32_t));
    num_groups = grouper->num_groups();
    if (n_aggs == 0 && out_keys) {
      auto uniq = grouper->GetUniques();
      if (!uniq.ok()) return Fail(uniq.status());
      for (int k = 0; k < n_keys; ++k) CopyOut(*uniq.ValueOrDie().values[k].array(), &out_keys[k]);
    }
  }
  if (out_num_groups) *out_num_groups = num_groups;
  return 0;
}

// compute::internal::GroupBy over CHUNKED inputs, hash_sum aggregates only: the one multi-batch form the reference
// computes correctly (multi-batch hash_count / hash_min_max are broken upstream; counts are obtained as
// hash_sum over an int64 column of ones carrying the value's validity).  keys[k * n_chunks + c] / args[j * n_chunks + c]
// are chunk c of key k / argument j; every chunk is its own buffer at offset 0 and at most 2^27 rows, so that
// ExecBatchIterator yields one batch per chunk and the int32 key-bytes limit of GrouperImpl::Consume is kept.
// This is how the reference sums a column longer than one batch: sums[g] += v in row order across the batches.
B200_API int ref_groupby_sum_chunked(int n_keys, int n_chunks, const b200_col* keys, int n_aggs, const b200_col* args,
                                     b200_out_col* out_aggs, b200_out_col* out_keys, uint32_t* out_num_groups) {
  compute::ExecContext ctx;
  auto chunked = [&](const b200_col* cols) {
    ArrayVector pieces;
    for (int c = 0; c < n_chunks; ++c) pieces.push_back(MakeArray(Wrap(cols[c])));
    return Datum(std::make_shared<ChunkedArray>(pieces, TypeOf(cols[0].type)));
  };
  std::vector<Datum> key_datums, arg_datums;
  for (int k = 0; k < n_keys; ++k) key_datums.push_back(chunked(keys + size_t(k) * n_chunks));
  std::vector<compute::internal::Aggregate> specs;
  for (int j = 0; j < n_aggs; ++j) {
    arg_datums.push_back(chunked(args + size_t(j) * n_chunks));
    specs.push_back({"hash_sum", nullptr});
  }
  auto res = compute::internal::GroupBy(arg_datums, key_datums, specs, &ctx);
  if (!res.ok()) return Fail(res.status());
  auto st = res.ValueOrDie().array();
  for (int j = 0; j < n_aggs; ++j) CopyOut(*st->child_data[j], &out_aggs[j]);
  for (int k = 0; k < n_keys; ++k) CopyOut(*st->child_data[n_aggs + k], &out_keys[k]);
  if (out_num_groups) *out_num_groups = static_cast<uint32_t>(st->length);
  return 0;
}

// Grouper::MakeGroupings (hash_aggregate.cc:968-1001): list<int32> offsets and values
B200_API int ref_make_groupings(const uint32_t* ids, int64_t n, uint32_t num_groups, int32_t* out_offsets, int32_t* out_indices) {
  compute::ExecContext ctx;
  auto buf = std::make_shared<Buffer>(reinterpret_cast<const uint8_t*>(ids), n * 4);
  UInt32Array arr(n, buf);
  auto res = compute::internal::Grouper::MakeGroupings(arr, num_groups, &ctx);
  if (!res.ok()) return Fail(res.status());
  auto list = res.ValueOrDie();
  std::memcpy(out_offsets, list->raw_value_offsets(), (size_t(num_groups) + 1) * 4);
  auto values = std::static_pointer_cast<Int32Array>(list->values());
  if (n > 0) std::memcpy(out_indices, values->raw_values(), size_t(n) * 4);
  return 0;
}

// "unique" / "value_counts" (compute/kernels/vector_hash.cc:723-782) on one fixed-width column.
// out_values: the distinct values in first-occurrence order (a null, if any, is one entry); out_counts nullable.
B200_API int ref_value_counts(const b200_col* values, b200_out_col* out_values, int64_t* out_counts, int64_t* out_n) {
  compute::ExecContext ctx;
  if (out_counts) {
    auto res = compute::CallFunction("value_counts", {Datum(Wrap(*values))}, &ctx);
    if (!res.ok()) return Fail(res.status());
    auto st = res.ValueOrDie().array();
    CopyOut(*st->child_data[0], out_values);
    std::memcpy(out_counts, st->child_data[1]->GetValues<int64_t>(1), size_t(st->length) * 8);
    *out_n = st->length;
    return 0;
  }
  auto res = compute::CallFunction("unique", {Datum(Wrap(*values))}, &ctx);
  if (!res.ok()) return Fail(res.status());
  auto arr = res.ValueOrDie().array();
  CopyOut(*arr, out_values);
  *out_n = arr->length;
  return 0;
}

// "dictionary_encode" (compute/kernels/vector_hash.cc:165-226, 648-668) on one fixed-width column.
// encode_nulls == 0: DictionaryEncodeOptions::MASK (null rows -> null indices, no null in the dictionary); != 0: ENCODE.
// out_indices: int32 per row; out_index_valid: one byte per row; out_dict: the dictionary; out_n: its length.
B200_API int ref_dictionary_encode(const b200_col* values, int encode_nulls, int32_t* out_indices, uint8_t* out_index_valid,
                                   b200_out_col* out_dict, int64_t* out_n) {
  compute::ExecContext ctx;
  compute::DictionaryEncodeOptions opts(encode_nulls ? compute::DictionaryEncodeOptions::ENCODE
                                                     : compute::DictionaryEncodeOptions::MASK);
  auto res = compute::CallFunction("dictionary_encode", {Datum(Wrap(*values))}, &opts, &ctx);
  if (!res.ok()) return Fail(res.status());
  auto arr = res.ValueOrDie().array();
  const int32_t* idx = arr->GetValues<int32_t>(1);
  for (int64_t i = 0; i < arr->length; ++i) {
    const bool valid = !arr->buffers[0] || BitUtil::GetBit(arr->buffers[0]->data(), arr->offset + i);
    out_indices[i] = valid ? idx[i] : 0;
    out_index_valid[i] = valid ? 1 : 0;
  }
  CopyOut(*arr->dictionary, out_dict);
  *out_n = arr->dictionary->length;
  return 0;
}

}  // extern "C"
