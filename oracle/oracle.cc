// TEST INFRASTRUCTURE ONLY — the CPU oracle.  Never linked, imported or executed by the product
// path (arrow1_b200/, include/); only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline
// / --impl reference legs may load it.
//
// A sequential restatement, in plain C++, of the reference's algorithm for the hot path
// (Arrow C++ 5.0.0-SNAPSHOT, paths under /root/reference/cpp/src/arrow/compute/kernels/):
//   oracle_filter        PrimitiveFilter + PrimitiveFilterImpl::Exec   vector_selection.cc:570-835
//                        GetFilterOutputSize                           vector_selection.cc:61-88
//   oracle_sort_indices  ArraySortIndices + ArrayCompareSorter         vector_sort.cc:374-405, 559-575
//                        PartitionNulls (stable, NaN before null)      vector_sort.cc:224-302
//   oracle_groupby       GrouperImpl::Consume (first-occurrence ids)   hash_aggregate.cc:358-408
//                        GroupedCountImpl / SumImpl / MinMaxImpl       hash_aggregate.cc:466-768
//                        hash_mean := MeanImpl::Finalize               aggregate_basic_internal.h:99-109
// Parity pinned: tests/test_oracle.py checks this file against the golden vectors transcribed
// from the reference's own tests (tests/golden/*.json) and, when oracle/_ref is built, against
// the unmodified reference (oracle/ref_driver.cc) on randomised inputs.
// Column descriptors reuse b200_col / b200_out_col (include/arrow1_b200.h) with HOST pointers.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <numeric>
#include <string>
#include <unordered_map>
#include <vector>

#include "arrow1_b200.h"

namespace {

inline bool GetBit(const void* bits, int64_t i) {
  return (static_cast<const uint8_t*>(bits)[i >> 3] >> (i & 7)) & 1;
}
inline void SetBitTo(void* bits, int64_t i, bool v) {
  uint8_t* b = static_cast<uint8_t*>(bits) + (i >> 3);
  uint8_t m = uint8_t(1u << (i & 7));
  *b = v ? (*b | m) : (*b & ~m);
}
inline int BitsOf(int t) {
  switch (t) {
    case B200_BOOL: return 1;
    case B200_UINT8: case B200_INT8: return 8;
    case B200_UINT16: case B200_INT16: return 16;
    case B200_UINT32: case B200_INT32: case B200_FLOAT: return 32;
    case B200_UINT64: case B200_INT64: case B200_DOUBLE: return 64;
  }
  return 0;
}
inline bool IsValid(const b200_col& c, int64_t i) { return c.validity == nullptr || GetBit(c.validity, c.offset + i); }

template <typename T>
inline T Value(const b200_col& c, int64_t i) {
  T v;
  std::memcpy(&v, static_cast<const char*>(c.data) + (c.offset + i) * sizeof(T), sizeof(T));
  return v;
}

// ---- filter -----------------------------------------------------------------------------
// Row semantics distilled from the block cases of PrimitiveFilterImpl::Exec
// (vector_selection.cc:605-730): a row is emitted as a VALUE when the filter slot is valid and
// true (validity bit = the value's validity, data = source bytes even under a null,
// WriteValue :733-735); under EMIT_NULL a null filter slot emits a NULL (validity 0, data
// zeroed, WriteNull :743-746).
int FilterImpl(const b200_col* values, const b200_col* mask, int null_sel, b200_out_col* out, int* has_validity) {
  if (mask->type != B200_BOOL) return B200_NOT_IMPLEMENTED;   // vector_selection.cc:1885-1887
  if (values->length != mask->length) return B200_INVALID;    // compute/exec.cc:186-189
  const int bits = BitsOf(values->type);
  if (bits == 0) return B200_NOT_IMPLEMENTED;
  // GetPrimitiveArg: null_count = 0 without a bitmap (util_internal.cc:56)
  const int64_t vnc = values->validity ? values->null_count : 0;
  const int64_t fnc = mask->validity ? mask->null_count : 0;
  // vector_selection.cc:799-812
  out->null_count = (vnc == 0 && (null_sel == 0 || fnc == 0)) ? 0 : -1;
  const bool allocate_validity = vnc != 0 || fnc != 0;
  if (has_validity) *has_validity = allocate_validity;
  const int64_t n = values->length;
  const int bw = bits / 8;
  int64_t o = 0;
  char* od = static_cast<char*>(out->data);
  for (int64_t i = 0; i < n; ++i) {
    const bool fvalid = mask->validity == nullptr || GetBit(mask->validity, mask->offset + i);
    const bool fbit = GetBit(mask->data, mask->offset + i);
    if (fvalid && fbit) {
      if (bits == 1) {
        SetBitTo(od, o, GetBit(values->data, values->offset + i));
      } else {
        std::memcpy(od + o * bw, static_cast<const char*>(values->data) + (values->offset + i) * bw, bw);
      }
      if (allocate_validity && out->validity) SetBitTo(out->validity, o, IsValid(*values, i));
      ++o;
    } else if (!fvalid && null_sel == 1) {
      if (bits == 1) {
        SetBitTo(od, o, false);
      } else {
        std::memset(od + o * bw, 0, bw);
      }
      if (out->validity) SetBitTo(out->validity, o, false);
      ++o;
    }
  }
  out->length = o;
  return B200_OK;
}

// ---- sort -------------------------------------------------------------------------------
template <typename T>
int SortImpl(const b200_col* c, int order, uint64_t base, uint64_t* out) {
  const int64_t n = c->length;
  std::iota(out, out + n, uint64_t(0));                                        // vector_sort.cc:568
  // PartitionNullsOnly then PartitionNullLikes, both std::stable_partition (vector_sort.cc:224-302)
  uint64_t* nulls_begin = out + n;
  if (c->validity) nulls_begin = std::stable_partition(out, out + n, [&](uint64_t i) { return IsValid(*c, int64_t(i)); });
  uint64_t* nan_begin = nulls_begin;
  if (std::is_floating_point<T>::value) {
    nan_begin = std::stable_partition(out, nulls_begin, [&](uint64_t i) {
      T v = Value<T>(*c, int64_t(i));
      return !(v != v);
    });
  }
  // ArrayCompareSorter: std::stable_sort with lhs < rhs / rhs < lhs (vector_sort.cc:386-402).
  // (The integer counting-sort path, :408-525, is also stable and therefore result-identical.)
  if (order == 0) {
    std::stable_sort(out, nan_begin, [&](uint64_t l, uint64_t r) { return Value<T>(*c, int64_t(l)) < Value<T>(*c, int64_t(r)); });
  } else {
    std::stable_sort(out, nan_begin, [&](uint64_t l, uint64_t r) { return Value<T>(*c, int64_t(r)) < Value<T>(*c, int64_t(l)); });
  }
  if (base) for (int64_t i = 0; i < n; ++i) out[i] += base;
  return B200_OK;
}

// ---- several sort keys: MultipleKeyComparator (vector_sort.cc:1050-1181) under one std::stable_sort.
// The reference's sorters (RadixRecordBatchSorter :955-1048, MultipleKeyRecordBatchSorter :1184-1356,
// MultipleKeyTableSorter :1381-1560) partition nulls / NaNs of the first key first, but every step is a stable
// sort or stable partition under this one ordering, so the result is the stable sort by the comparator.
template <typename T>
int CompareKey(const b200_col& c, int order, uint64_t l, uint64_t r) {
  const bool lv = IsValid(c, int64_t(l)), rv = IsValid(c, int64_t(r));
  if (!lv && !rv) return 0;          // :1108-1116 null handling
  if (!lv) return 1;
  if (!rv) return -1;
  const T a = Value<T>(c, int64_t(l)), b = Value<T>(c, int64_t(r));
  if (std::is_floating_point<T>::value) {   // :1146-1156 NaN handling
    const bool an = a != a, bn = b != b;
    if (an && bn) return 0;
    if (an) return 1;
    if (bn) return -1;
  }
  int cmp = a == b ? 0 : (a > b ? 1 : -1);
  return order ? -cmp : cmp;
}

int CompareKeyDispatch(const b200_col& c, int order, uint64_t l, uint64_t r) {
  switch (c.type) {
    case B200_UINT8: return CompareKey<uint8_t>(c, order, l, r);
    case B200_INT8: return CompareKey<int8_t>(c, order, l, r);
    case B200_UINT16: return CompareKey<uint16_t>(c, order, l, r);
    case B200_INT16: return CompareKey<int16_t>(c, order, l, r);
    case B200_UINT32: return CompareKey<uint32_t>(c, order, l, r);
    case B200_INT32: return CompareKey<int32_t>(c, order, l, r);
    case B200_UINT64: return CompareKey<uint64_t>(c, order, l, r);
    case B200_INT64: return CompareKey<int64_t>(c, order, l, r);
    case B200_FLOAT: return CompareKey<float>(c, order, l, r);
    default: return CompareKey<double>(c, order, l, r);
  }
}

// ---- group-by ---------------------------------------------------------------------------
struct AggState {
  b200_agg_spec spec;
  // sum / mean / count
  std::vector<double> fsum;
  std::vector<uint64_t> isum;   // int64 and uint64 sums share two's-complement addition
  std::vector<int64_t> count;
  // min_max, kept in the three accumulate domains
  std::vector<double> fmin, fmax;
  std::vector<int64_t> imin, imax;
  std::vector<uint64_t> umin, umax;
  std::vector<uint8_t> has_values, has_nulls;
};

inline bool IsFloat(int t) { return t == B200_FLOAT || t == B200_DOUBLE; }
inline bool IsSigned(int t) { return t == B200_INT8 || t == B200_INT16 || t == B200_INT32 || t == B200_INT64; }

inline double LoadF(const b200_col& c, int64_t i) { return c.type == B200_FLOAT ? double(Value<float>(c, i)) : Value<double>(c, i); }
inline int64_t LoadI(const b200_col& c, int64_t i) {
  switch (c.type) {
    case B200_INT8: return Value<int8_t>(c, i);
    case B200_INT16: return Value<int16_t>(c, i);
    case B200_INT32: return Value<int32_t>(c, i);
    default: return Value<int64_t>(c, i);
  }
}
inline uint64_t LoadU(const b200_col& c, int64_t i) {
  switch (c.type) {
    case B200_UINT8: return Value<uint8_t>(c, i);
    case B200_UINT16: return Value<uint16_t>(c, i);
    case B200_UINT32: return Value<uint32_t>(c, i);
    default: return Value<uint64_t>(c, i);
  }
}

void StoreTyped(int type, void* dst, int64_t g, double f, int64_t i, uint64_t u) {
  switch (type) {
    case B200_FLOAT: static_cast<float*>(dst)[g] = float(f); break;
    case B200_DOUBLE: static_cast<double*>(dst)[g] = f; break;
    case B200_INT8: static_cast<int8_t*>(dst)[g] = int8_t(i); break;
    case B200_INT16: static_cast<int16_t*>(dst)[g] = int16_t(i); break;
    case B200_INT32: static_cast<int32_t*>(dst)[g] = int32_t(i); break;
    case B200_INT64: static_cast<int64_t*>(dst)[g] = i; break;
    case B200_UINT8: static_cast<uint8_t*>(dst)[g] = uint8_t(u); break;
    case B200_UINT16: static_cast<uint16_t*>(dst)[g] = uint16_t(u); break;
    case B200_UINT32: static_cast<uint32_t*>(dst)[g] = uint32_t(u); break;
    case B200_UINT64: static_cast<uint64_t*>(dst)[g] = u; break;
  }
}

}  // namespace

extern "C" {

B200_API const char* oracle_version() { return "oracle port of arrow-5.0.0-SNAPSHOT filter/sort_indices/hash_aggregate"; }

B200_API int oracle_filter(const b200_col* values, const b200_col* mask, int null_selection, b200_out_col* out,
                           int* has_validity) {
  return FilterImpl(values, mask, null_selection, out, has_validity);
}

// PrimitiveTake / PrimitiveTakeImpl / BooleanTakeImpl (vector_selection.cc:273-469, 493-526) with CheckIndexBounds
// (util/int_util.cc): out[i] = values[indices[i]]; a null index or a null value gives a null with zero data; the
// validity bitmap is always produced; null_count exact.  Returns 7 (IndexError) on an out-of-bounds index.
B200_API int oracle_take(const b200_col* values, const b200_col* indices, int boundscheck, b200_out_col* out) {
  const int vbits = BitsOf(values->type);
  const int ibits = BitsOf(indices->type);
  const bool isigned = indices->type == B200_INT8 || indices->type == B200_INT16 || indices->type == B200_INT32 ||
                       indices->type == B200_INT64;
  if (ibits < 8 || indices->type == B200_FLOAT || indices->type == B200_DOUBLE) return 3;   // TypeError
  const uint8_t* ivalid = static_cast<const uint8_t*>(indices->validity);
  const uint8_t* vvalid = static_cast<const uint8_t*>(values->validity);
  const int64_t n = indices->length;
  auto get_bit = [](const uint8_t* b, int64_t i) { return (b[i >> 3] >> (i & 7)) & 1; };
  auto index_at = [&](int64_t i, bool* negative) -> uint64_t {
    const int64_t p = indices->offset + i;
    *negative = false;
    switch (indices->type) {
      case B200_INT8: { int8_t v = static_cast<const int8_t*>(indices->data)[p]; *negative = v < 0; return uint64_t(int64_t(v)); }
      case B200_UINT8: return static_cast<const uint8_t*>(indices->data)[p];
      case B200_INT16: { int16_t v = static_cast<const int16_t*>(indices->data)[p]; *negative = v < 0; return uint64_t(int64_t(v)); }
      case B200_UINT16: return static_cast<const uint16_t*>(indices->data)[p];
      case B200_INT32: { int32_t v = static_cast<const int32_t*>(indices->data)[p]; *negative = v < 0; return uint64_t(int64_t(v)); }
      case B200_UINT32: return static_cast<const uint32_t*>(indices->data)[p];
      case B200_INT64: { int64_t v = static_cast<const int64_t*>(indices->data)[p]; *negative = v < 0; return uint64_t(v); }
      default: return static_cast<const uint64_t*>(indices->data)[p];
    }
  };
  (void)isigned;
  if (boundscheck) {
    for (int64_t i = 0; i < n; ++i) {
      if (ivalid && !get_bit(ivalid, indices->offset + i)) continue;
      bool neg;
      const uint64_t ix = index_at(i, &neg);
      if (neg || ix >= uint64_t(values->length)) return 7;
    }
  }
  uint8_t* od = static_cast<uint8_t*>(out->data);
  uint8_t* ov = static_cast<uint8_t*>(out->validity);
  std::memset(ov, 0, size_t((n + 7) / 8));
  if (vbits == 1) std::memset(od, 0, size_t((n + 7) / 8));
  int64_t valid_count = 0;
  for (int64_t i = 0; i < n; ++i) {
    bool ok = !ivalid || get_bit(ivalid, indices->offset + i);
    uint64_t ix = 0;
    if (ok) {
      bool neg;
      ix = index_at(i, &neg);
      ok = !vvalid || get_bit(vvalid, values->offset + int64_t(ix));
    }
    if (ok) {
      ++valid_count;
      ov[i >> 3] |= uint8_t(1u << (i & 7));
      const int64_t src = values->offset + int64_t(ix);
      if (vbits == 1) {
        if (get_bit(static_cast<const uint8_t*>(values->data), src)) od[i >> 3] |= uint8_t(1u << (i & 7));
      } else {
        std::memcpy(od + i * (vbits / 8), static_cast<const uint8_t*>(values->data) + src * (vbits / 8), size_t(vbits / 8));
      }
    } else if (vbits != 1) {
      std::memset(od + i * (vbits / 8), 0, size_t(vbits / 8));
    }
  }
  out->length = n;
  out->null_count = n - valid_count;
  return 0;
}

B200_API int oracle_filter_count(const b200_col* mask, int null_sel, int64_t* out_len) {
  // GetFilterOutputSize (vector_selection.cc:61-88)
  int64_t n = 0;
  for (int64_t i = 0; i < mask->length; ++i) {
    const bool fvalid = mask->validity == nullptr || GetBit(mask->validity, mask->offset + i);
    const bool fbit = GetBit(mask->data, mask->offset + i);
    n += null_sel ? (fbit || !fvalid) : (fbit && fvalid);
  }
  *out_len = n;
  return B200_OK;
}

B200_API int oracle_sort_indices(const b200_col* values, int order, uint64_t index_base, uint64_t* out) {
  switch (values->type) {
    case B200_UINT8: return SortImpl<uint8_t>(values, order, index_base, out);
    case B200_INT8: return SortImpl<int8_t>(values, order, index_base, out);
    case B200_UINT16: return SortImpl<uint16_t>(values, order, index_base, out);
    case B200_INT16: return SortImpl<int16_t>(values, order, index_base, out);
    case B200_UINT32: return SortImpl<uint32_t>(values, order, index_base, out);
    case B200_INT32: return SortImpl<int32_t>(values, order, index_base, out);
    case B200_UINT64: return SortImpl<uint64_t>(values, order, index_base, out);
    case B200_INT64: return SortImpl<int64_t>(values, order, index_base, out);
    case B200_FLOAT: return SortImpl<float>(values, order, index_base, out);
    case B200_DOUBLE: return SortImpl<double>(values, order, index_base, out);
  }
  return B200_NOT_IMPLEMENTED;  // boolean has no array_sort_indices kernel (vector_sort.cc:577-606)
}

B200_API int oracle_sort_indices_multi(const b200_col* keys, const int* orders, int n_keys, uint64_t* out) {
  if (n_keys < 1) return B200_INVALID;   // "Must specify one or more sort keys" (vector_sort.cc:1636)
  const int64_t n = keys[0].length;
  for (int k = 0; k < n_keys; ++k) {
    if (BitsOf(keys[k].type) < 8) return B200_NOT_IMPLEMENTED;
    if (keys[k].length != n) return B200_INVALID;
  }
  std::iota(out, out + n, uint64_t(0));
  std::stable_sort(out, out + n, [&](uint64_t l, uint64_t r) {
    for (int k = 0; k < n_keys; ++k) {
      const int c = CompareKeyDispatch(keys[k], orders[k], l, r);
      if (c != 0) return c < 0;
    }
    return false;
  });
  return B200_OK;
}

B200_API int oracle_groupby(int n_keys, const b200_col* keys, int n_aggs, const b200_agg_spec* aggs,
                            const b200_col* args, b200_out_col* out_aggs, b200_out_col* out_keys,
                            uint32_t* out_num_groups, uint32_t* out_ids) {
  if (n_keys <= 0) return B200_INVALID;
  const int64_t n = keys[0].length;
  for (int k = 0; k < n_keys; ++k) {
    if (BitsOf(keys[k].type) == 0) return B200_NOT_IMPLEMENTED;
    if (keys[k].length != n) return B200_INVALID;
  }
  for (int j = 0; j < n_aggs; ++j) {
    if (args[j].length != n) return B200_INVALID;
    const int t = args[j].type;
    // GroupedSumImpl / GroupedMinMaxImpl reject boolean (hash_aggregate.cc:550-556, 684-694)
    if ((aggs[j].kind == B200_HASH_SUM || aggs[j].kind == B200_HASH_MEAN || aggs[j].kind == B200_HASH_MIN_MAX) &&
        (t == B200_BOOL || BitsOf(t) == 0))
      return B200_NOT_IMPLEMENTED;
  }
  // KeyEncoder: per key column [null byte][raw bytes, zero if null] (hash_aggregate.cc:48-196);
  // boolean keys encode one byte 0/1 (:97-140)
  std::vector<int> widths(n_keys);
  int row_bytes = 0;
  for (int k = 0; k < n_keys; ++k) {
    widths[k] = keys[k].type == B200_BOOL ? 1 : BitsOf(keys[k].type) / 8;
    row_bytes += 1 + widths[k];
  }
  std::unordered_map<std::string, uint32_t> map;
  std::vector<std::string> uniques;
  std::vector<AggState> st(n_aggs);
  for (int j = 0; j < n_aggs; ++j) st[j].spec = aggs[j];
  std::string enc(row_bytes, '\0');
  for (int64_t i = 0; i < n; ++i) {
    char* p = &enc[0];
    for (int k = 0; k < n_keys; ++k) {
      const b200_col& c = keys[k];
      if (!IsValid(c, i)) {
        *p++ = 1;
        std::memset(p, 0, widths[k]);
      } else {
        *p++ = 0;
        if (c.type == B200_BOOL) {
          *p = GetBit(c.data, c.offset + i) ? 1 : 0;
        } else {
          std::memcpy(p, static_cast<const char*>(c.data) + (c.offset + i) * widths[k], widths[k]);
        }
      }
      p += widths[k];
    }
    // GrouperImpl::Consume: new key => id = num_groups_++ (hash_aggregate.cc:391-403)
    auto it = map.emplace(enc, uint32_t(uniques.size()));
    const uint32_t g = it.first->second;
    if (it.second) {
      uniques.push_back(enc);
      for (auto& a : st) {
        // MaybeReserve with zero / anti-extreme fill (hash_aggregate.cc:450-458, 627-640, 708-718)
        a.fsum.push_back(0); a.isum.push_back(0); a.count.push_back(0);
        a.fmin.push_back(std::numeric_limits<double>::infinity());
        a.fmax.push_back(-std::numeric_limits<double>::infinity());
        a.imin.push_back(std::numeric_limits<int64_t>::max());
        a.imax.push_back(std::numeric_limits<int64_t>::min());
        a.umin.push_back(std::numeric_limits<uint64_t>::max());
        a.umax.push_back(0);
        a.has_values.push_back(0); a.has_nulls.push_back(0);
      }
    }
    if (out_ids) out_ids[i] = g;
    for (int j = 0; j < n_aggs; ++j) {
      AggState& a = st[j];
      const b200_col& c = args[j];
      const bool valid = IsValid(c, i);
      switch (a.spec.kind) {
        case B200_HASH_COUNT:  // hash_aggregate.cc:474-505
          a.count[g] += a.spec.option ? !valid : valid;
          break;
        case B200_HASH_SUM:
        case B200_HASH_MEAN:   // sums[g] += v; counts[g] += 1, row order (hash_aggregate.cc:533-545)
          if (valid) {
            if (IsFloat(c.type)) a.fsum[g] += LoadF(c, i);
            else if (IsSigned(c.type)) a.isum[g] += uint64_t(LoadI(c, i));
            else a.isum[g] += LoadU(c, i);
            a.count[g] += 1;
          }
          break;
        case B200_HASH_MIN_MAX:  // std::max / std::min, has_values / has_nulls (hash_aggregate.cc:663-677)
          if (valid) {
            if (c.type == B200_FLOAT) {
              // compare in float like the reference (CType = float); widening is exact
              float v = Value<float>(c, i);
              a.fmax[g] = double(std::max(float(a.fmax[g]), v));
              a.fmin[g] = double(std::min(float(a.fmin[g]), v));
            } else if (c.type == B200_DOUBLE) {
              double v = Value<double>(c, i);
              a.fmax[g] = std::max(a.fmax[g], v);
              a.fmin[g] = std::min(a.fmin[g], v);
            } else if (IsSigned(c.type)) {
              int64_t v = LoadI(c, i);
              a.imax[g] = std::max(a.imax[g], v);
              a.imin[g] = std::min(a.imin[g], v);
            } else {
              uint64_t v = LoadU(c, i);
              a.umax[g] = std::max(a.umax[g], v);
              a.umin[g] = std::min(a.umin[g], v);
            }
            a.has_values[g] = 1;
          } else {
            a.has_nulls[g] = 1;
          }
          break;
      }
    }
  }
  const uint32_t G = uint32_t(uniques.size());
  if (out_num_groups) *out_num_groups = G;
  // finalize (hash_aggregate.cc:507-517, 591-612, 722-757)
  int o = 0;
  for (int j = 0; j < n_aggs && out_aggs; ++j) {
    AggState& a = st[j];
    const int t = args[j].type;
    switch (a.spec.kind) {
      case B200_HASH_COUNT: {
        b200_out_col& oc = out_aggs[o++];
        oc.length = G; oc.null_count = 0;
        for (uint32_t g = 0; g < G; ++g) static_cast<int64_t*>(oc.data)[g] = a.count[g];
        if (oc.validity) { std::memset(oc.validity, 0, (G + 7) / 8); for (uint32_t g = 0; g < G; ++g) SetBitTo(oc.validity, g, true); }
        break;
      }
      case B200_HASH_SUM: {
        b200_out_col& oc = out_aggs[o++];
        oc.length = G; oc.null_count = 0;
        if (oc.validity) std::memset(oc.validity, 0, (G + 7) / 8);
        for (uint32_t g = 0; g < G; ++g) {
          if (IsFloat(t)) static_cast<double*>(oc.data)[g] = a.fsum[g];
          else static_cast<uint64_t*>(oc.data)[g] = a.isum[g];
          const bool valid = a.count[g] > 0;
          oc.null_count += !valid;
          if (oc.validity) SetBitTo(oc.validity, g, valid);
        }
        break;
      }
      case B200_HASH_MEAN: {
        b200_out_col& oc = out_aggs[o++];
        oc.length = G; oc.null_count = 0;
        if (oc.validity) std::memset(oc.validity, 0, (G + 7) / 8);
        for (uint32_t g = 0; g < G; ++g) {
          const bool valid = a.count[g] > 0;
          double s = IsFloat(t) ? a.fsum[g] : (IsSigned(t) ? double(int64_t(a.isum[g])) : double(a.isum[g]));
          // MeanImpl::Finalize: static_cast<double>(sum) / count (aggregate_basic_internal.h:104-106)
          static_cast<double*>(oc.data)[g] = valid ? s / double(a.count[g]) : 0.0;
          oc.null_count += !valid;
          if (oc.validity) SetBitTo(oc.validity, g, valid);
        }
        break;
      }
      case B200_HASH_MIN_MAX: {
        b200_out_col& omin = out_aggs[o++];
        b200_out_col& omax = out_aggs[o++];
        omin.length = omax.length = G;
        omin.null_count = 0;
        if (omin.validity) std::memset(omin.validity, 0, (G + 7) / 8);
        if (omax.validity) std::memset(omax.validity, 0, (G + 7) / 8);
        for (uint32_t g = 0; g < G; ++g) {
          StoreTyped(t, omin.data, g, a.fmin[g], a.imin[g], a.umin[g]);
          StoreTyped(t, omax.data, g, a.fmax[g], a.imax[g], a.umax[g]);
          bool valid = a.has_values[g];
          if (a.spec.option == 1) valid = valid && !a.has_nulls[g];  // EMIT_NULL: BitmapAndNot (:727-732)
          omin.null_count += !valid;
          if (omin.validity) SetBitTo(omin.validity, g, valid);
          if (omax.validity) SetBitTo(omax.validity, g, valid);
        }
        omax.null_count = omin.null_count;
        break;
      }
    }
  }
  // GetUniques / Decode (hash_aggregate.cc:412-429, 65-94, 174-192)
  for (int k = 0; k < n_keys && out_keys; ++k) {
    b200_out_col& oc = out_keys[k];
    oc.length = G; oc.null_count = 0;
    int off = 0;
    for (int kk = 0; kk < k; ++kk) off += 1 + widths[kk];
    if (oc.validity) std::memset(oc.validity, 0, (G + 7) / 8);
    if (keys[k].type == B200_BOOL && oc.data) std::memset(oc.data, 0, (G + 7) / 8);
    for (uint32_t g = 0; g < G; ++g) {
      const char* p = uniques[g].data() + off;
      const bool valid = p[0] == 0;
      oc.null_count += !valid;
      if (oc.validity) SetBitTo(oc.validity, g, valid);
      if (oc.data) {
        if (keys[k].type == B200_BOOL) SetBitTo(oc.data, g, p[1] != 0);
        else std::memcpy(static_cast<char*>(oc.data) + size_t(g) * widths[k], p + 1, widths[k]);
      }
    }
  }
  return B200_OK;
}

}  // extern "C"
