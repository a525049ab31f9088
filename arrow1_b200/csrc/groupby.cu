// Grouper + grouped aggregates: open-addressing hash aggregation.
// Replaces GrouperImpl::Consume / GetUniques, GroupedCountImpl, GroupedSumImpl, GroupedMinMaxImpl
// and the GroupBy driver (reference compute/kernels/hash_aggregate.cc:313-437, 466-768, 890-966);
// hash_mean is MeanImpl::Finalize (aggregate_basic_internal.h:99-109) over the sum state.
//
//   key      the reference's row-major encoding [null byte][raw bytes, zero if null] per key column
//            (hash_aggregate.cc:48-196) is never materialised: the same information is packed into
//            one 64-bit register key = value bits | one null flag per column (floats by bit
//            pattern, so -0.0 != 0.0 and NaNs group by payload, hash_aggregate_test.cc:280-289).
//   table    slots of `stride` bytes {key, first_row, per-aggregate state}; linear probing; keys
//            claimed with 64-bit atomicCAS.  Two extra slots hold the null key and the key equal to
//            the EMPTY sentinel when a single 64-bit key column leaves no spare bit.
//   order    group ids are first-occurrence order (hash_aggregate.cc:391-403): every slot keeps
//            min(row_base + i); finalize ranks the occupied slots by it with the radix sort.
//   kernels  gb_consume_kernel<false>  rows -> global table, global atomics (high cardinality)
//            gb_consume_kernel<true>   rows -> per-CTA shared-memory table (shared atomics), flushed
//                                      once per CTA into the global table (low cardinality)
//            gb_rehash / gb_merge / gb_collect / gb_emit / gb_dense_consume
// Algorithmic HBM bytes per row: every key and argument column read once:
// C3 12.125 B/row, C4 16 B/row.
#include <stddef.h>

#include <algorithm>
#include <vector>

#include "common.cuh"

// kernel-variant switches of the timing experiments (profiles/): environment lookups only in a debug build
#ifdef B200_DEBUG_KNOBS
#include <stdlib.h>
#define B200_KNOB(name) getenv(name)
#else
#define B200_KNOB(name) (static_cast<const char*>(nullptr))
#endif

namespace {

constexpr int kMaxKeys = 4;
constexpr int kMaxCols = 8;
constexpr int kMaxAggs = 12;
constexpr uint64_t kEmpty = ~uint64_t(0);
constexpr uint32_t kNoId = 0xFFFFFFFFu;
constexpr int kMaxProbe = 256;
constexpr int kLocalProbe = 24;

enum Domain { kSigned = 0, kUnsigned = 1, kFloat = 2 };
enum Mode { kModeFused = 0, kModeInsert = 1, kModeLookup = 2 };

struct GbKey {
  const void* data;  // shifted by the element offset (non-bool)
  BitmapRef bits;    // bool key data
  BitmapRef vv;
  int32_t type, nbits, shift, pad;
};
struct GbCol {
  const void* data;  // shifted
  BitmapRef bits;    // bool argument (hash_count only)
  BitmapRef vv;
  int32_t type, pad;
};
struct GbAgg {
  int32_t kind, dom, option, off, col, arg_type;
};
struct GbSchema {
  GbAgg aggs[kMaxAggs];
  int32_t n_aggs;
  int32_t n_keys;
  int32_t null_shift;   // bit position of the first null flag, or -1 (single 64-bit key)
  int32_t stride;       // global slot stride (bytes, multiple of 16)
  int32_t lstride;      // shared slot stride (bytes, multiple of 8)
  int32_t key_bits[kMaxKeys], key_shift[kMaxKeys], key_type[kMaxKeys];
};
struct GbTable {
  char* slots;       // (cap + 2) slots
  uint64_t cap;      // power of two
};
struct GbArgs {
  GbSchema s;
  GbKey keys[kMaxKeys];
  GbCol cols[kMaxCols];
  int32_t n_cols;
  int32_t mode;
  int64_t n;
  uint64_t row_base;
  const uint64_t* row_ids;   // nullable: explicit global row number of every row (repartitioned input)
  GbTable t;
  uint32_t* counters;        // [0] groups, [1] deferred rows, [2] claimed slots (grouper), [3] error
  const uint32_t* defer_in;  // nullable: only rows whose bit is set are processed
  const uint32_t* select;    // nullable: caller's selection bitmap (bit offset 0); rows whose bit is clear do not exist
  uint32_t* defer_out;       // rows that found no slot
  uint32_t* claimed;         // grouper: newly claimed slot indices
  uint32_t* slot_ids;        // grouper: dense id per slot
  uint32_t* out_ids;         // grouper lookup output
  char* spill;               // fast path: slot images of groups the global table had no room for (counters[5])
  uint64_t* zf_scratch;      // fast path: per-CTA earliest-zero-row words [grid][kFastGroups]
  int32_t dbg;               // B200_GB_DBG: phase-skipping bits for timing experiments (results are wrong)
  int32_t local_slots;       // shared-memory table size (power of two), 0 = none
  int32_t rows_per_block;
};

__device__ __forceinline__ uint64_t mix64(uint64_t k) {
  k ^= k >> 33;
  k *= 0xff51afd7ed558ccdull;
  k ^= k >> 33;
  k *= 0xc4ceb9fe1a85ec53ull;
  k ^= k >> 33;
  return k;
}

// ---- order-preserving encodings for min/max -------------------------------------------------
__device__ __forceinline__ uint64_t ord_f64(double v) {  // v is not NaN; +-0 canonicalised
  uint64_t b = __double_as_longlong(v);
  if ((b << 1) == 0) b = 0;
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double unord_f64(uint64_t o) {
  uint64_t b = (o >> 63) ? (o & 0x7FFFFFFFFFFFFFFFull) : ~o;
  return __longlong_as_double(b);
}

struct Val {
  uint64_t bits;  // signed: sign-extended, unsigned: zero-extended, float: f64 bits (f32 widened)
  bool valid;
};

__device__ __forceinline__ Val load_val(const GbCol& c, int64_t row) {
  Val v;
  v.valid = bitmap_get(c.vv, row);
  v.bits = 0;
  switch (c.type) {
    case B200_BOOL: v.bits = bitmap_get(c.bits, row); break;
    case B200_UINT8: v.bits = reinterpret_cast<const uint8_t*>(c.data)[row]; break;
    case B200_INT8: v.bits = uint64_t(int64_t(reinterpret_cast<const int8_t*>(c.data)[row])); break;
    case B200_UINT16: v.bits = reinterpret_cast<const uint16_t*>(c.data)[row]; break;
    case B200_INT16: v.bits = uint64_t(int64_t(reinterpret_cast<const int16_t*>(c.data)[row])); break;
    case B200_UINT32: v.bits = reinterpret_cast<const uint32_t*>(c.data)[row]; break;
    case B200_INT32: v.bits = uint64_t(int64_t(reinterpret_cast<const int32_t*>(c.data)[row])); break;
    case B200_FLOAT: v.bits = __double_as_longlong(double(reinterpret_cast<const float*>(c.data)[row])); break;
    default: v.bits = reinterpret_cast<const uint64_t*>(c.data)[row]; break;  // 64-bit types
  }
  return v;
}

// packed key of a row; *special = 1 (null single-64-bit key) or 2 (key == EMPTY sentinel)
__device__ __forceinline__ uint64_t load_key(const GbArgs& a, int64_t row, int* special) {
  uint64_t key = 0;
  *special = 0;
#pragma unroll
  for (int k = 0; k < kMaxKeys; ++k) {
    if (k < a.s.n_keys) {
      const GbKey& c = a.keys[k];
      const bool valid = bitmap_get(c.vv, row);
      if (!valid) {
        if (a.s.null_shift >= 0) key |= uint64_t(1) << (a.s.null_shift + k);
        else *special = 1;
        continue;
      }
      uint64_t v;
      switch (c.nbits) {
        case 1: v = bitmap_get(c.bits, row); break;
        case 8: v = reinterpret_cast<const uint8_t*>(c.data)[row]; break;
        case 16: v = reinterpret_cast<const uint16_t*>(c.data)[row]; break;
        case 32: v = reinterpret_cast<const uint32_t*>(c.data)[row]; break;
        default: v = reinterpret_cast<const uint64_t*>(c.data)[row]; break;
      }
      key |= v << c.shift;
    }
  }
  if (a.s.null_shift < 0 && *special == 0 && key == kEmpty) *special = 2;
  return key;
}

// ---- slot state -----------------------------------------------------------------------------
// COUNT: {u64 count}  SUM/MEAN: {u64|f64 sum, u64 count}
// MIN_MAX: {u64 min_ord, u64 max_ord, u64 zero_first, u64 flags(bit0 has_values, bit1 has_nulls)}
__host__ __device__ inline int state_bytes(int kind) {
  return kind == B200_HASH_COUNT ? 8 : (kind == B200_HASH_MIN_MAX ? 32 : 16);
}

__device__ __forceinline__ void init_slot(char* slot, const GbSchema& s) {
  uint64_t* w = reinterpret_cast<uint64_t*>(slot);
  w[0] = kEmpty;
  w[1] = ~uint64_t(0);
  for (int j = 0; j < s.n_aggs; ++j) {
    uint64_t* st = reinterpret_cast<uint64_t*>(slot + s.aggs[j].off);
    if (s.aggs[j].kind == B200_HASH_MIN_MAX) {
      st[0] = ~uint64_t(0);
      st[1] = 0;
      st[2] = ~uint64_t(0);
      st[3] = 0;
    } else {
      st[0] = 0;
      if (s.aggs[j].kind != B200_HASH_COUNT) st[1] = 0;
    }
  }
}

__device__ __forceinline__ void atomic_min_checked(uint64_t* p, uint64_t v) {
  // a stale read is always >= the true value (it only decreases), so skipping is safe
  if (v < *reinterpret_cast<volatile uint64_t*>(p)) atomicMin(reinterpret_cast<unsigned long long*>(p), (unsigned long long)v);
}
__device__ __forceinline__ void atomic_max_checked(uint64_t* p, uint64_t v) {
  if (v > *reinterpret_cast<volatile uint64_t*>(p)) atomicMax(reinterpret_cast<unsigned long long*>(p), (unsigned long long)v);
}
__device__ __forceinline__ void atomic_or_checked(uint64_t* p, uint64_t v) {
  if ((*reinterpret_cast<volatile uint64_t*>(p) & v) != v) atomicOr(reinterpret_cast<unsigned long long*>(p), (unsigned long long)v);
}

// one row into one aggregate's state
__device__ __forceinline__ void accumulate(char* slot, const GbAgg& g, const Val& v, uint64_t row) {
  uint64_t* st = reinterpret_cast<uint64_t*>(slot + g.off);
  switch (g.kind) {
    case B200_HASH_COUNT:  // hash_aggregate.cc:474-505
      if (g.option ? !v.valid : v.valid) atomicAdd(reinterpret_cast<unsigned long long*>(st), 1ull);
      break;
    case B200_HASH_SUM:
    case B200_HASH_MEAN:   // hash_aggregate.cc:533-545
      if (v.valid) {
        if (g.dom == kFloat) atomicAdd(reinterpret_cast<double*>(st), __longlong_as_double(v.bits));
        else atomicAdd(reinterpret_cast<unsigned long long*>(st), (unsigned long long)v.bits);
        atomicAdd(reinterpret_cast<unsigned long long*>(st + 1), 1ull);
      }
      break;
    default:               // hash_aggregate.cc:663-677
      if (!v.valid) {
        atomic_or_checked(st + 3, 2);
        break;
      }
      atomic_or_checked(st + 3, 1);
      uint64_t o;
      if (g.dom == kFloat) {
        const double d = __longlong_as_double(v.bits);
        if (d != d) break;  // std::max/min never pick a NaN, but has_values is set
        // +-0.0 ties: the earlier row's sign wins (std::max(cur, v) keeps cur when equal)
        if (d == 0.0) atomic_min_checked(st + 2, (row << 1) | (v.bits >> 63));
        o = ord_f64(d);
      } else if (g.dom == kSigned) {
        o = v.bits ^ 0x8000000000000000ull;
      } else {
        o = v.bits;
      }
      atomic_min_checked(st + 0, o);
      atomic_max_checked(st + 1, o);
      break;
  }
}

// fold a source slot's state (plain memory) into a destination slot (atomics)
__device__ __forceinline__ void merge_state(char* dst, const char* src, const GbSchema& s) {
  const uint64_t* sw = reinterpret_cast<const uint64_t*>(src);
  atomic_min_checked(reinterpret_cast<uint64_t*>(dst) + 1, sw[1]);
  for (int j = 0; j < s.n_aggs; ++j) {
    const GbAgg& g = s.aggs[j];
    uint64_t* d = reinterpret_cast<uint64_t*>(dst + g.off);
    const uint64_t* c = reinterpret_cast<const uint64_t*>(src + g.off);
    switch (g.kind) {
      case B200_HASH_COUNT:
        if (c[0]) atomicAdd(reinterpret_cast<unsigned long long*>(d), (unsigned long long)c[0]);
        break;
      case B200_HASH_SUM:
      case B200_HASH_MEAN:
        if (c[1]) {
          if (g.dom == kFloat) atomicAdd(reinterpret_cast<double*>(d), __longlong_as_double(c[0]));
          else atomicAdd(reinterpret_cast<unsigned long long*>(d), (unsigned long long)c[0]);
          atomicAdd(reinterpret_cast<unsigned long long*>(d + 1), (unsigned long long)c[1]);
        }
        break;
      default:
        atomic_min_checked(d + 0, c[0]);
        atomic_max_checked(d + 1, c[1]);
        atomic_min_checked(d + 2, c[2]);
        if (c[3]) atomic_or_checked(d + 3, c[3]);
        break;
    }
  }
}

// find or claim the global slot of `key`; nullptr when the probe limit is hit (table too full)
__device__ __forceinline__ char* global_slot(const GbTable& t, int stride, uint64_t key, int special,
                                             bool insert, bool* claimed_new, uint64_t* slot_index) {
  *claimed_new = false;
  if (special) {
    const uint64_t idx = t.cap + (special - 1);
    char* slot = t.slots + idx * stride;
    uint64_t* kp = reinterpret_cast<uint64_t*>(slot);
    *slot_index = idx;
    // occupancy marker of a special slot: key word 0 (EMPTY = unused)
    if (*reinterpret_cast<volatile uint64_t*>(kp) == kEmpty) {
      if (!insert) return nullptr;
      if (atomicCAS(reinterpret_cast<unsigned long long*>(kp), (unsigned long long)kEmpty, 0ull) == kEmpty) *claimed_new = true;
    }
    return slot;
  }
  uint64_t idx = mix64(key) & (t.cap - 1);
  for (int probe = 0; probe < kMaxProbe; ++probe) {
    char* slot = t.slots + idx * stride;
    uint64_t* kp = reinterpret_cast<uint64_t*>(slot);
    uint64_t cur = *reinterpret_cast<volatile uint64_t*>(kp);
    if (cur == kEmpty) {
      if (!insert) return nullptr;
      cur = atomicCAS(reinterpret_cast<unsigned long long*>(kp), (unsigned long long)kEmpty, (unsigned long long)key);
      if (cur == kEmpty) {
        *claimed_new = true;
        *slot_index = idx;
        return slot;
      }
    }
    if (cur == key) {
      *slot_index = idx;
      return slot;
    }
    idx = (idx + 1) & (t.cap - 1);
  }
  return nullptr;
}

// one thread per 8-byte word: consecutive threads write consecutive words (the per-slot form wrote 8 bytes at a
// stride of one slot per store instruction: 1.8 TB/s)
__global__ void gb_init_kernel(const __grid_constant__ GbSchema s, char* slots, uint64_t n_slots, int stride) {
  __shared__ uint64_t image[64];
  const int words = stride / 8;
  if (threadIdx.x < 64) image[threadIdx.x] = 0;
  __syncthreads();
  if (threadIdx.x == 0) init_slot(reinterpret_cast<char*>(image), s);
  __syncthreads();
  const uint64_t total = n_slots * uint64_t(words);
  uint64_t* out = reinterpret_cast<uint64_t*>(slots);
  const uint64_t step = uint64_t(gridDim.x) * blockDim.x;
  uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  int w = int(i % uint64_t(words));
  const int step_w = int(step % uint64_t(words));
  for (; i < total; i += step) {
    out[i] = image[w];
    w += step_w;
    if (w >= words) w -= words;
  }
}

#include "groupby_hp.cuh"

// ---- the consume kernel ---------------------------------------------------------------------
template <bool kLocal>
__global__ void __launch_bounds__(kLocal ? 1024 : 256) gb_consume_kernel(const __grid_constant__ GbArgs a) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ uint32_t s_new_groups;
  __shared__ uint32_t s_deferred;
  char* ltab = reinterpret_cast<char*>(smem);
  const int lstride = a.s.lstride;
  const uint32_t lmask = uint32_t(a.local_slots - 1);
  if (threadIdx.x == 0) {
    s_new_groups = 0;
    s_deferred = 0;
  }
  if (kLocal) {
    for (int i = threadIdx.x; i < a.local_slots; i += blockDim.x) init_slot(ltab + i * lstride, a.s);
  }
  __syncthreads();

  const int64_t begin = int64_t(blockIdx.x) * a.rows_per_block;
  const int64_t end = min(a.n, begin + a.rows_per_block);
  // trip count made warp-uniform so that the warp can be re-converged every iteration (lanes that
  // diverge in probe / CAS loops otherwise keep running apart: 2-3 active lanes per instruction)
  for (int64_t row0 = begin + (threadIdx.x & ~31); row0 < end; row0 += blockDim.x) {
    __syncwarp();
    const int64_t row = row0 + (threadIdx.x & 31);
    if (row >= end) continue;
    if (a.defer_in && !((a.defer_in[row >> 5] >> (row & 31)) & 1u)) continue;
    if (a.select && !((a.select[row >> 5] >> (row & 31)) & 1u)) continue;
    int special;
    const uint64_t key = load_key(a, row, &special);
    const uint64_t grow = a.row_ids ? a.row_ids[row] : a.row_base + uint64_t(row);
    char* slot = nullptr;
    bool local_hit = false;
    if (kLocal && !special) {
      uint32_t idx = uint32_t(mix64(key) >> 20) & lmask;
      for (int probe = 0; probe < kLocalProbe; ++probe) {
        char* ls = ltab + idx * lstride;
        uint64_t* kp = reinterpret_cast<uint64_t*>(ls);
        uint64_t cur = *reinterpret_cast<volatile uint64_t*>(kp);
        if (cur == kEmpty) cur = atomicCAS(reinterpret_cast<unsigned long long*>(kp), (unsigned long long)kEmpty, (unsigned long long)key);
        if (cur == kEmpty || cur == key) {
          slot = ls;
          local_hit = true;
          break;
        }
        idx = (idx + 1) & lmask;
      }
    }
    uint64_t slot_index = 0;
    if (!local_hit) {
      bool claimed;
      slot = global_slot(a.t, a.s.stride, key, special, a.mode != kModeLookup, &claimed, &slot_index);
      if (claimed) {
        atomicAdd(&s_new_groups, 1u);
        if (a.claimed) a.claimed[atomicAdd(a.counters + 2, 1u)] = uint32_t(slot_index);
      }
      if (!slot) {
        if (a.mode == kModeLookup) {
          atomicExch(a.counters + 3, 1u);  // every key was inserted by the pass before: cannot happen
        } else {
          atomicOr(a.defer_out + (row >> 5), 1u << (row & 31));
          atomicAdd(&s_deferred, 1u);
        }
        continue;
      }
    }
    if (a.mode == kModeLookup) {
      a.out_ids[row] = a.slot_ids[slot_index];
      continue;
    }
    atomic_min_checked(reinterpret_cast<uint64_t*>(slot) + 1, grow);
    if (a.mode == kModeInsert) continue;
    // each distinct argument column is loaded once and feeds all of its aggregates
#pragma unroll
    for (int c = 0; c < kMaxCols; ++c) {
      if (c < a.n_cols) {
        const Val v = load_val(a.cols[c], row);
        for (int j = 0; j < a.s.n_aggs; ++j)
          if (a.s.aggs[j].col == c) accumulate(slot, a.s.aggs[j], v, grow);
      }
    }
  }
  __syncthreads();
  if (kLocal) {
    // flush the CTA's partial aggregates; the host sized the global table so this cannot fail
    for (int i = threadIdx.x; i < a.local_slots; i += blockDim.x) {
      const char* ls = ltab + i * lstride;
      const uint64_t key = *reinterpret_cast<const uint64_t*>(ls);
      if (key == kEmpty) continue;
      bool claimed;
      uint64_t slot_index;
      char* gs = global_slot(a.t, a.s.stride, key, 0, true, &claimed, &slot_index);
      if (!gs) {
        // no room left in the global table: the group goes to the spill list as a slot image;
        // the host grows the table and imports the list
        if (a.mode == kModeInsert) {
          atomicExch(a.counters + 3, 2u);
          continue;
        }
        uint64_t* img = reinterpret_cast<uint64_t*>(a.spill + size_t(atomicAdd(a.counters + 5, 1u)) * a.s.stride);
        const uint64_t* lw = reinterpret_cast<const uint64_t*>(ls);
        for (int w = 0; w < lstride / 8; ++w) img[w] = lw[w];
        img[a.s.stride / 8 - 1] = 0;
        continue;
      }
      if (claimed) {
        atomicAdd(&s_new_groups, 1u);
        if (a.claimed) a.claimed[atomicAdd(a.counters + 2, 1u)] = uint32_t(slot_index);
      }
      if (a.mode == kModeInsert) atomic_min_checked(reinterpret_cast<uint64_t*>(gs) + 1, reinterpret_cast<const uint64_t*>(ls)[1]);
      else merge_state(gs, ls, a.s);
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    if (s_new_groups) atomicAdd(a.counters + 0, s_new_groups);
    if (s_deferred) atomicAdd(a.counters + 1, s_deferred);
  }
}

// ---- fast low-cardinality path ----------------------------------------------------------------
// One value column feeding any mix of hash_sum / hash_mean / hash_count / hash_min_max (C3: 12.125
// algorithmic B/row), up to 1024 groups per CTA.  Shared-memory atomics on sm_100a are native only
// for 32-bit integers (measured 6.5 updates/cycle/SM; 64-bit and floating-point ones are CAS loops
// at 0.07-1.3, and their retry loops split the warp), so nothing is accumulated with atomics:
// every 4096-row tile is counting-sorted by group in shared memory and reduced per group.
//   P1  128-bit loads (4 consecutive rows per thread, two sub-tiles); key -> dense group id through
//       a shared hash table (new keys are inserted in a separate, rarely taken step); the row's rank
//       inside its group = native ATOMS.ADD on the group's tile counter
//   P2  exclusive scan of the 1024 tile counters
//   P3  values scattered to sorted[offset[group] + rank]
//   P4  each thread owns two groups: walks their segments with plain adds / min / max in registers
//       and updates the CTA's per-group accumulators without any conflict
// Rows that cannot be placed (special keys, more than 1024 groups in a CTA) are flagged in the
// deferred-row bitmap and taken by the generic kernel afterwards.  At the end every CTA folds its
// groups into the global table.
constexpr int kFastThreads = 512;
constexpr int kFastSub = 4;                              // consecutive rows per thread per sub-tile
constexpr int kFastSubs = 2;                             // sub-tiles per tile
constexpr int kFastRows = kFastSub * kFastSubs;          // rows per thread per tile
constexpr int kFastSubRows = kFastThreads * kFastSub;    // 2048
constexpr int kFastTile = kFastThreads * kFastRows;      // 4096
constexpr int kFastGroups = 1024;                        // dense group ids per CTA
constexpr int kFastMaxGroupsHint = 1008;                 // cardinality hint up to which the path is chosen
constexpr uint32_t kGidUnset = 0xFFFFu, kGidOverflow = 0xFFFEu;
constexpr uint32_t kNoRank = 0xFFFFFFFFu;

// Hash table of a CTA: 16 KB of keys in 16-byte buckets read with one LDS.128 — four 32-bit keys
// (KEY32: a single key column of at most 32 bits without nulls; 4096 slots) or two 64-bit packed
// keys (2048 slots) — plus a 16-bit dense group id per slot.  With <= 1024 groups a key sits in its
// home bucket with probability > 0.99 (four-key buckets), so a warp rarely probes twice.
constexpr int kFastKeyBytes = 16 * 1024;
constexpr int kFastSlots32 = kFastKeyBytes / 4;   // 4096
constexpr int kFastSlots64 = kFastKeyBytes / 8;   // 2048
constexpr int kFastBuckets = kFastKeyBytes / 16;  // 1024
constexpr uint32_t kEmpty32 = 0xFFFFFFFFu;

struct FastSmem {
  uint4 hbuckets[kFastBuckets];
  uint64_t sorted[kFastTile];
  uint64_t sum[kFastGroups];
  uint64_t mn[kFastGroups];
  uint64_t mx[kFastGroups];
  uint32_t tcnt[kFastGroups];
  uint32_t toff[kFastGroups];
  uint32_t cnt[kFastGroups];
  uint32_t nulls[kFastGroups];
  uint32_t first[kFastGroups];
  uint16_t hgid[kFastSlots32];
  uint32_t warp_tot[kFastThreads / 32];
  uint32_t n_groups, need_insert, new_groups, deferred;
};

__device__ __forceinline__ uint32_t fast_hash(uint64_t key) {
  uint32_t x = uint32_t(key) ^ (uint32_t(key >> 32) * 0x9E3779B9u);
  x *= 0x85EBCA6Bu;
  x ^= x >> 15;
  x *= 0xC2B2AE35u;
  return x >> (32 - 10);   // bucket index
}
static_assert(kFastBuckets == 1024, "fast_hash yields 10 bits");

// slot of `key` in bucket b: KEY32 -> 4 candidates, else 2; -1 if absent; *has_empty = bucket has room
template <bool KEY32, class SM>
__device__ __forceinline__ int fast_match(const SM& S, uint32_t b, uint64_t key, bool* has_empty) {
  uint4 w;
  asm volatile("ld.volatile.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(w.x), "=r"(w.y), "=r"(w.z), "=r"(w.w)
               : "r"(uint32_t(__cvta_generic_to_shared(&S.hbuckets[b]))));
  if (KEY32) {
    const uint32_t k = uint32_t(key);
    *has_empty = w.w == kEmpty32;   // slots of a bucket fill in order
    const uint32_t m = (w.x == k ? 1u : 0u) | (w.y == k ? 2u : 0u) | (w.z == k ? 4u : 0u) | (w.w == k ? 8u : 0u);
    return m ? int(4 * b) + (__ffs(m) - 1) : -1;
  } else {
    const uint64_t k0 = (uint64_t(w.y) << 32) | w.x, k1 = (uint64_t(w.w) << 32) | w.z;
    *has_empty = k1 == kEmpty;
    const uint32_t m = (k0 == key ? 1u : 0u) | (k1 == key ? 2u : 0u);
    return m ? int(2 * b) + (__ffs(m) - 1) : -1;
  }
}

template <class SM>
__device__ __forceinline__ int fast_gid_of_slot(const SM& S, int slot) {
  const uint32_t g = *reinterpret_cast<const volatile uint16_t*>(&S.hgid[slot]);
  return g < uint32_t(kFastGroups) ? int(g) : (g == kGidUnset ? -1 : -2);
}

// the rare continuation of a lookup whose home bucket neither holds the key nor has room
template <bool KEY32, class SM>
__device__ __noinline__ int fast_lookup_far(const SM& S, uint64_t key, uint32_t b) {
#pragma unroll 1
  for (int probe = 1; probe < kLocalProbe; ++probe) {
    b = (b + 1) & (kFastBuckets - 1);
    bool has_empty;
    const int slot = fast_match<KEY32>(S, b, key, &has_empty);
    if (slot >= 0) return fast_gid_of_slot(S, slot);
    if (has_empty) return -1;
  }
  return -2;
}

// group id of `key`, -1 = not (yet) in the table, -2 = cannot be placed
template <bool KEY32, class SM>
__device__ __forceinline__ int fast_lookup(const SM& S, uint64_t key) {
  const uint32_t b = fast_hash(key);
  bool has_empty;
  const int slot = fast_match<KEY32>(S, b, key, &has_empty);
  if (slot >= 0) return fast_gid_of_slot(S, slot);
  if (has_empty) return -1;
  return fast_lookup_far<KEY32>(S, key, b);
}

template <bool KEY32, class SM>
__device__ __forceinline__ void fast_insert(SM& S, uint64_t key) {
  uint32_t b = fast_hash(key);
#pragma unroll 1
  for (int probe = 0; probe < kLocalProbe; ++probe) {
    bool has_empty;
    if (fast_match<KEY32>(S, b, key, &has_empty) >= 0) return;
    if (has_empty) {
      if (*reinterpret_cast<const volatile uint32_t*>(&S.n_groups) >= uint32_t(kFastGroups)) return;
      // claim the first free slot of the bucket; a lost race re-examines the same bucket
      const int per = KEY32 ? 4 : 2;
      bool mine = false, same = false;
      int slot = -1;
      for (int i = 0; i < per && !mine && !same; ++i) {
        slot = int(per * b) + i;
        if (KEY32) {
          uint32_t* kp = reinterpret_cast<uint32_t*>(S.hbuckets) + slot;
          const uint32_t old = atomicCAS(kp, kEmpty32, uint32_t(key));
          mine = old == kEmpty32;
          same = old == uint32_t(key);
        } else {
          unsigned long long* kp = reinterpret_cast<unsigned long long*>(S.hbuckets) + slot;
          const unsigned long long old = atomicCAS(kp, (unsigned long long)kEmpty, (unsigned long long)key);
          mine = old == kEmpty;
          same = old == key;
        }
      }
      if (same) return;
      if (mine) {
        const uint32_t g = atomicAdd(&S.n_groups, 1u);
        S.hgid[slot] = uint16_t(g < uint32_t(kFastGroups) ? g : kGidOverflow);
        return;
      }
      // the bucket filled up under us: fall through to the next one
    }
    b = (b + 1) & (kFastBuckets - 1);
  }
}

template <int KMODE, bool VFAST, bool DIRECT>
__global__ void __launch_bounds__(kFastThreads, 2) gb_fast_kernel(const __grid_constant__ GbArgs a) {
  extern __shared__ __align__(16) unsigned char smem[];
  FastSmem& S = *reinterpret_cast<FastSmem*>(smem);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  for (int i = t; i < kFastBuckets; i += kFastThreads) S.hbuckets[i] = make_uint4(kEmpty32, kEmpty32, kEmpty32, kEmpty32);
  for (int i = t; i < kFastSlots32; i += kFastThreads) S.hgid[i] = uint16_t(kGidUnset);
  for (int i = t; i < kFastGroups; i += kFastThreads) {
    S.sum[i] = 0;
    S.mn[i] = ~uint64_t(0);
    S.mx[i] = 0;
    S.tcnt[i] = 0;
    S.cnt[i] = 0;
    S.nulls[i] = 0;
    S.first[i] = 0xFFFFFFFFu;
  }
  if (t == 0) {
    S.n_groups = 0;
    S.need_insert = 0;
    S.new_groups = 0;
    S.deferred = 0;
  }
  __syncthreads();
  // this launch serves the aggregates whose argument is column 0 (col == 0); a multi-column group-by runs
  // the kernel once per distinct argument column, the others carry col == -1
  int dom = kSigned;
  bool want_minmax = false;
  for (int j = a.s.n_aggs - 1; j >= 0; --j)
    if (a.s.aggs[j].col == 0) {
      dom = a.s.aggs[j].dom;
      if (a.s.aggs[j].kind == B200_HASH_MIN_MAX) want_minmax = true;
    }
  const int64_t begin = int64_t(blockIdx.x) * a.rows_per_block;
  const int64_t end = min(a.n, begin + a.rows_per_block);   // a.n is a multiple of kFastTile here
  uint64_t* zf = a.zf_scratch + size_t(blockIdx.x) * kFastGroups;

#pragma unroll 1
  for (int64_t base = begin; base < end; base += kFastTile) {
    uint64_t vbits[kFastRows];
    int gid[kFastRows];
    uint32_t vmask = 0;   // validity of the values
    uint32_t newkey = 0;  // rows whose key was not in the table when the tile started
    // ---- P1a: loads and group lookup
    {
      uint64_t key[kFastRows];
      uint32_t special = 0;
#pragma unroll
      for (int s = 0; s < kFastSubs; ++s) {
        const int64_t r0 = base + s * kFastSubRows + t * kFastSub;
        if (KMODE == 1) {
          const uint4 k4 = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint32_t*>(a.keys[0].data) + r0);
          key[4 * s + 0] = k4.x; key[4 * s + 1] = k4.y; key[4 * s + 2] = k4.z; key[4 * s + 3] = k4.w;
#pragma unroll
          for (int i = 0; i < kFastSub; ++i)
            if (uint32_t(key[4 * s + i]) == kEmpty32) special |= 1u << (4 * s + i);
        } else if (KMODE == 2) {
          const ulonglong2* kp = reinterpret_cast<const ulonglong2*>(reinterpret_cast<const uint64_t*>(a.keys[0].data) + r0);
          const ulonglong2 ka = kp[0], kb = kp[1];
          key[4 * s + 0] = ka.x; key[4 * s + 1] = ka.y; key[4 * s + 2] = kb.x; key[4 * s + 3] = kb.y;
#pragma unroll
          for (int i = 0; i < kFastSub; ++i)
            if (key[4 * s + i] == kEmpty) special |= 1u << (4 * s + i);
        } else {
#pragma unroll
          for (int i = 0; i < kFastSub; ++i) {
            int sp;
            key[4 * s + i] = load_key(a, r0 + i, &sp);
            if (sp) special |= 1u << (4 * s + i);
          }
        }
        if (VFAST) {
          const ulonglong2* vp = reinterpret_cast<const ulonglong2*>(reinterpret_cast<const uint64_t*>(a.cols[0].data) + r0);
          const ulonglong2 va = vp[0], vb = vp[1];
          vbits[4 * s + 0] = va.x; vbits[4 * s + 1] = va.y; vbits[4 * s + 2] = vb.x; vbits[4 * s + 3] = vb.y;
          uint32_t m = 0xFu;
          if (a.cols[0].vv.words) {
            const int64_t bit = a.cols[0].vv.bit_offset + r0;
            const uint8_t* bytes = reinterpret_cast<const uint8_t*>(a.cols[0].vv.words);
            const int sh = int(bit & 7);
            m = uint32_t(__ldg(bytes + (bit >> 3))) >> sh;
            if (sh > 4) m |= uint32_t(__ldg(bytes + (bit >> 3) + 1)) << (8 - sh);
          }
          vmask |= (m & 0xFu) << (4 * s);
        } else {
#pragma unroll
          for (int i = 0; i < kFastSub; ++i) {
            const Val v = load_val(a.cols[0], r0 + i);
            vbits[4 * s + i] = v.bits;
            vmask |= (v.valid ? 1u : 0u) << (4 * s + i);
          }
        }
      }
      uint32_t pending = 0;
#pragma unroll
      for (int i = 0; i < kFastRows; ++i) {
        gid[i] = ((special >> i) & 1u) ? -2 : fast_lookup<KMODE == 1>(S, key[i]);
        if (gid[i] == -1) pending |= 1u << i;
      }
      newkey = pending;
      if (__any_sync(0xffffffffu, pending != 0) && lane == 0) S.need_insert = 1;
      __syncthreads();
      if (S.need_insert) {   // CTA-uniform; taken for the first tile(s) only
#pragma unroll
        for (int i = 0; i < kFastRows; ++i)
          if ((pending >> i) & 1u) fast_insert<KMODE == 1>(S, key[i]);
        __syncthreads();
        if (t == 0) S.need_insert = 0;
#pragma unroll
        for (int i = 0; i < kFastRows; ++i)
          if ((pending >> i) & 1u) {
            gid[i] = fast_lookup<KMODE == 1>(S, key[i]);
            if (gid[i] == -1) gid[i] = -2;
          }
        __syncwarp();
      }
    }
    if (DIRECT) {
      // ---- direct accumulation: native 32-bit atomics for counts, warp-converged CAS retries for
      // float sums (two rows in flight), read-then-rarely-CAS for the extremes
      uint32_t ok = 0;
#pragma unroll
      for (int i = 0; i < kFastRows; ++i)
        if (gid[i] >= 0) ok |= 1u << i;
      const uint32_t live = vmask & ok, nullm = ok & ~vmask, nk = newkey & ok, defer = ~ok & ((1u << kFastRows) - 1);
#pragma unroll
      for (int i = 0; i < kFastRows; ++i)
        if ((live >> i) & 1u) atomicAdd(&S.cnt[gid[i]], 1u);
      if (nullm) {
#pragma unroll
        for (int i = 0; i < kFastRows; ++i)
          if ((nullm >> i) & 1u) atomicAdd(&S.nulls[gid[i]], 1u);
      }
      if (nk) {
#pragma unroll
        for (int i = 0; i < kFastRows; ++i)
          if ((nk >> i) & 1u) {
            const int64_t row = base + (i / kFastSub) * kFastSubRows + t * kFastSub + (i % kFastSub);
            atomicMin(&S.first[gid[i]], uint32_t(row - begin));
          }
      }
      if (defer) {
#pragma unroll
        for (int i = 0; i < kFastRows; ++i)
          if ((defer >> i) & 1u) {
            const int64_t row = base + (i / kFastSub) * kFastSubRows + t * kFastSub + (i % kFastSub);
            atomicOr(a.defer_out + (row >> 5), 1u << (row & 31));
          }
        atomicAdd(&S.deferred, uint32_t(__popc(defer)));
      }
      __syncwarp();
      if (dom == kFloat) {
#pragma unroll
        for (int i = 0; i < kFastRows; i += 2) {
          uint32_t pend = (live >> i) & 3u;
          while (__any_sync(0xffffffffu, pend != 0)) {
            uint64_t o0 = 0, o1 = 0;
            if (pend & 1u) o0 = *reinterpret_cast<const volatile uint64_t*>(&S.sum[gid[i]]);
            if (pend & 2u) o1 = *reinterpret_cast<const volatile uint64_t*>(&S.sum[gid[i + 1]]);
            if (pend & 1u) {
              const uint64_t nw = __double_as_longlong(__longlong_as_double(o0) + __longlong_as_double(vbits[i]));
              if (atomicCAS(reinterpret_cast<unsigned long long*>(&S.sum[gid[i]]), (unsigned long long)o0, (unsigned long long)nw) == o0) pend &= ~1u;
            }
            if (pend & 2u) {
              const uint64_t nw = __double_as_longlong(__longlong_as_double(o1) + __longlong_as_double(vbits[i + 1]));
              if (atomicCAS(reinterpret_cast<unsigned long long*>(&S.sum[gid[i + 1]]), (unsigned long long)o1, (unsigned long long)nw) == o1) pend &= ~2u;
            }
          }
        }
      } else {
#pragma unroll
        for (int i = 0; i < kFastRows; ++i) {
          if (!((live >> i) & 1u)) continue;
          uint32_t* w = reinterpret_cast<uint32_t*>(&S.sum[gid[i]]);
          const uint32_t lo = uint32_t(vbits[i]), hi = uint32_t(vbits[i] >> 32);
          const uint32_t old_lo = atomicAdd(w, lo);
          const uint32_t carry = (old_lo + lo) < old_lo ? 1u : 0u;
          if (hi + carry) atomicAdd(w + 1, hi + carry);
        }
      }
      if (want_minmax) {
#pragma unroll
        for (int i = 0; i < kFastRows; ++i) {
          bool lt = false, gt = false;
          uint64_t o = 0;
          if ((live >> i) & 1u) {
            bool use = true;
            if (dom == kFloat) {
              const double dv = __longlong_as_double(vbits[i]);
              use = dv == dv;
              o = ord_f64(dv);
              if ((vbits[i] << 1) == 0) {
                const int64_t row = base + (i / kFastSub) * kFastSubRows + t * kFastSub + (i % kFastSub);
                atomic_min_checked(zf + gid[i], ((a.row_base + uint64_t(row)) << 1) | (vbits[i] >> 63));
              }
            } else {
              o = dom == kSigned ? (vbits[i] ^ 0x8000000000000000ull) : vbits[i];
            }
            if (use) {
              lt = o < *reinterpret_cast<const volatile uint64_t*>(&S.mn[gid[i]]);
              gt = o > *reinterpret_cast<const volatile uint64_t*>(&S.mx[gid[i]]);
            }
          }
          if (__any_sync(0xffffffffu, lt || gt)) {
            if (lt) atomicMin(reinterpret_cast<unsigned long long*>(&S.mn[gid[i]]), (unsigned long long)o);
            if (gt) atomicMax(reinterpret_cast<unsigned long long*>(&S.mx[gid[i]]), (unsigned long long)o);
            __syncwarp();
          }
        }
      }
      continue;
    }
    // ---- P1b: rank inside the group (native shared atomics), null counts, first rows, deferred rows
    uint32_t rank[kFastRows];
    uint32_t defer = 0;
#pragma unroll
    for (int i = 0; i < kFastRows; ++i) {
      rank[i] = kNoRank;
      const int64_t row = base + (i / kFastSub) * kFastSubRows + t * kFastSub + (i % kFastSub);
      if (gid[i] < 0) {
        defer |= 1u << i;
        continue;
      }
      // first rows: a key is inserted in the tile where the CTA first meets it, and every row of
      // that key in that tile was pending, so only those rows can lower the group's first row
      if ((newkey >> i) & 1u) atomicMin(&S.first[gid[i]], uint32_t(row - begin));
      if ((vmask >> i) & 1u) {
        rank[i] = atomicAdd(&S.tcnt[gid[i]], 1u);
        // +-0.0 ties in hash_min_max: the earliest zero row decides the sign (see accumulate())
        if (want_minmax && dom == kFloat && (vbits[i] << 1) == 0)
          atomic_min_checked(zf + gid[i], ((a.row_base + uint64_t(row)) << 1) | (vbits[i] >> 63));
      } else {
        atomicAdd(&S.nulls[gid[i]], 1u);
      }
    }
    if (defer) {
#pragma unroll
      for (int i = 0; i < kFastRows; ++i)
        if ((defer >> i) & 1u) {
          const int64_t row = base + (i / kFastSub) * kFastSubRows + t * kFastSub + (i % kFastSub);
          atomicOr(a.defer_out + (row >> 5), 1u << (row & 31));
        }
      atomicAdd(&S.deferred, uint32_t(__popc(defer)));
    }
    __syncthreads();
    // ---- P2: exclusive scan of the tile counters (two per thread)
    {
      const uint32_t c0 = S.tcnt[2 * t], c1 = S.tcnt[2 * t + 1];
      uint32_t incl = c0 + c1;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
      }
      if (lane == 31) S.warp_tot[warp] = incl;
      __syncthreads();
      uint32_t wpre = 0;
#pragma unroll
      for (int w = 0; w < kFastThreads / 32; ++w)
        if (w < warp) wpre += S.warp_tot[w];
      const uint32_t excl = wpre + incl - (c0 + c1);
      S.toff[2 * t] = excl;
      S.toff[2 * t + 1] = excl + c0;
    }
    __syncthreads();
    // ---- P3: scatter the values into group order
#pragma unroll
    for (int i = 0; i < kFastRows; ++i)
      if (rank[i] != kNoRank) S.sorted[S.toff[gid[i]] + rank[i]] = vbits[i];
    __syncthreads();
    // ---- P4: per-group reduction by the owning thread
#pragma unroll
    for (int q = 0; q < kFastGroups / kFastThreads; ++q) {
      const int g = t + q * kFastThreads;
      const uint32_t c = S.tcnt[g];
      if (c == 0) continue;
      const uint32_t o = S.toff[g];
      S.tcnt[g] = 0;
      S.cnt[g] += c;
      if (dom == kFloat) {
        double s = 0.0, lo = __longlong_as_double(0x7FF0000000000000ull), hi = __longlong_as_double(0xFFF0000000000000ull);
        const uint64_t* seg = S.sorted + o;
        uint32_t j = 0;
#pragma unroll 1
        for (; j + 2 <= c; j += 2) {
          const double v0 = __longlong_as_double(seg[j]), v1 = __longlong_as_double(seg[j + 1]);
          s += v0;
          s += v1;
          lo = v0 < lo ? v0 : lo;   // a NaN never compares true: it never wins (hash_aggregate.cc:672-673)
          hi = v0 > hi ? v0 : hi;
          lo = v1 < lo ? v1 : lo;
          hi = v1 > hi ? v1 : hi;
        }
        if (j < c) {
          const double v = __longlong_as_double(seg[j]);
          s += v;
          lo = v < lo ? v : lo;
          hi = v > hi ? v : hi;
        }
        S.sum[g] = __double_as_longlong(__longlong_as_double(S.sum[g]) + s);
        const uint64_t olo = ord_f64(lo), ohi = ord_f64(hi);
        if (olo < S.mn[g]) S.mn[g] = olo;
        if (ohi > S.mx[g]) S.mx[g] = ohi;
      } else {
        const uint64_t flip = dom == kSigned ? 0x8000000000000000ull : 0ull;
        uint64_t s = 0, lo = ~uint64_t(0), hi = 0;
#pragma unroll 1
        for (uint32_t j = 0; j < c; ++j) {
          const uint64_t v = S.sorted[o + j];
          s += v;
          const uint64_t ov = v ^ flip;
          lo = ov < lo ? ov : lo;
          hi = ov > hi ? ov : hi;
        }
        S.sum[g] += s;
        if (lo < S.mn[g]) S.mn[g] = lo;
        if (hi > S.mx[g]) S.mx[g] = hi;
      }
    }
    // no barrier here: the next tile's P1a touches neither tcnt nor sorted, and its own barrier
    // orders this P4 before the next P1b
  }
  __syncthreads();
  // ---- flush: every group of this CTA is folded into the global table
  constexpr bool KEY32 = KMODE == 1;
  for (int slot = t; slot < (KEY32 ? kFastSlots32 : kFastSlots64); slot += kFastThreads) {
    const uint32_t g = S.hgid[slot];
    if (g >= uint32_t(kFastGroups)) continue;   // unset or overflow
    const uint64_t key = KEY32 ? uint64_t(reinterpret_cast<const uint32_t*>(S.hbuckets)[slot])
                               : reinterpret_cast<const uint64_t*>(S.hbuckets)[slot];
    bool claimed;
    uint64_t slot_index;
    char* gs = global_slot(a.t, a.s.stride, key, 0, true, &claimed, &slot_index);
    if (!gs) {
      // no room (the table is sized from the cardinality hint, not for the worst case): the group
      // goes to the spill list as a fresh slot image; the host grows the table and imports it
      gs = a.spill + size_t(atomicAdd(a.counters + 5, 1u)) * a.s.stride;
      init_slot(gs, a.s);
      reinterpret_cast<uint64_t*>(gs)[0] = key;
      reinterpret_cast<uint64_t*>(gs)[a.s.stride / 8 - 1] = 0;
      claimed = false;
    }
    if (claimed) atomicAdd(&S.new_groups, 1u);
    const uint64_t cnt = S.cnt[g], nulls = S.nulls[g];
    if (S.first[g] != 0xFFFFFFFFu) atomic_min_checked(reinterpret_cast<uint64_t*>(gs) + 1, a.row_base + uint64_t(begin) + S.first[g]);
    for (int j = 0; j < a.s.n_aggs; ++j) {
      const GbAgg& ag = a.s.aggs[j];
      if (ag.col != 0) continue;
      uint64_t* d = reinterpret_cast<uint64_t*>(gs + ag.off);
      if (ag.kind == B200_HASH_COUNT) {
        const uint64_t c = ag.option ? nulls : cnt;
        if (c) atomicAdd(reinterpret_cast<unsigned long long*>(d), (unsigned long long)c);
      } else if (ag.kind == B200_HASH_MIN_MAX) {
        if (cnt) {
          atomic_min_checked(d + 0, S.mn[g]);
          atomic_max_checked(d + 1, S.mx[g]);
          atomic_min_checked(d + 2, zf[g]);
        }
        const uint64_t flags = (cnt ? 1ull : 0ull) | (nulls ? 2ull : 0ull);
        if (flags) atomic_or_checked(d + 3, flags);
      } else if (cnt) {
        if (ag.dom == kFloat) atomicAdd(reinterpret_cast<double*>(d), __longlong_as_double(S.sum[g]));
        else atomicAdd(reinterpret_cast<unsigned long long*>(d), (unsigned long long)S.sum[g]);
        atomicAdd(reinterpret_cast<unsigned long long*>(d + 1), (unsigned long long)cnt);
      }
    }
  }
  __syncthreads();
  if (t == 0) {
    if (S.new_groups) atomicAdd(a.counters + 0, S.new_groups);
    if (S.deferred) atomicAdd(a.counters + 1, S.deferred);
  }
}

// ---- binned low-cardinality path (v2 of the fast path) -------------------------------------------
// Same contract as gb_fast_kernel.  What changed, and why (profiles/r01_gb_v2_ncu_summary.txt: 4.9 warp
// instructions and 1.36 shared-memory wavefronts per row, both at their limit):
//   * one CTA of 1024 threads per SM and 8192-row tiles: a group sees ~8 rows per tile instead of ~4, so
//     the per-group bookkeeping of the reduction step is amortised over twice the rows
//   * every group owns a fixed-capacity bin of kBinCap values, laid out value-slot-major
//     (bins[j * 1024 + g]): a row's slot is known as soon as its rank comes back from the native
//     ATOMS.ADD, so the tile-wide scan of the counters, the per-row offset lookup and the registers that
//     held keys / values / ranks across two barriers are gone; and the owner threads of a warp read
//     consecutive addresses (conflict-free LDS.64) instead of 32 unrelated segments
//   * rows that find their bin full (skewed keys) are accumulated on the spot with the warp-converged
//     shared-memory CAS path, so skew costs speed but never defers rows
constexpr int kBinThreads = 1024;
constexpr int kBinRows = kFastSub * kFastSubs;          // rows per thread per tile (8)
constexpr int kBinSubRows = kBinThreads * kFastSub;     // 4096
constexpr int kBinTile = kBinThreads * kBinRows;        // 8192
// Two knobs measured on C3 (1 B rows, 1 K keys, B200) and left at their defaults (scripts/build_variants.sh):
//   B200_BIN_CAP  16 -> 14 puts BinSmem under 196 KB, i.e. 60 KB of L1 instead of 28 KB: 6.36 -> 6.51 ms (slower: more
//                 rows on the CAS overflow path, and the larger L1 buys nothing)
//   B200_BIN_LD   1 = column loads with L1::no_allocate, 2 = the four values of a thread as ONE 256-bit load: 6.36 ms
//                 either way.  Together with the steady-state form of the row loop (-25 % instructions, -2 % time)
//                 these pointed away from the memory system — wrongly: the ncu source page of that form showed 20 %
//                 of the samples on two register moves placed right behind the column loads (see the loop head of
//                 gb_bin32_kernel); with one load site per buffer and the L2 prefetch below the kernel is at 5.84 ms
#ifndef B200_BIN_CAP
#define B200_BIN_CAP 16
#endif
#ifndef B200_BIN_LD
#define B200_BIN_LD 0
#endif
// tiles whose key / value bytes are prefetched into L2 ahead of the demand loads, see gb_bin32_kernel.  Measured on C3
// (kernel ms per 1 B rows): 0 (off) 6.47, 1 5.84, 2 5.91, 3 6.08, 4 6.21
#ifndef B200_BIN_PF
#define B200_BIN_PF 1
#endif
constexpr int kBinCap = B200_BIN_CAP;                   // values per group per tile before the overflow path
constexpr int kBinSlots = 4096;                         // KEY32 table: {key, group id} pairs, load <= 1/4
constexpr uint32_t kBinGidOverflow = 0xFFFFFFFEu;     // == -2 as int: "cannot be placed"

struct BinSmem {
  // key -> dense group id.  KEY32 (one <= 32-bit key column without nulls): 4096 {key, gid} pairs probed
  // with ONE LDS.64 — the profile of the bucketed table (LDS.128 + four compares + a second load for
  // the id) showed 37 % of the kernel's instructions in the lookup.  Other key shapes keep the
  // bucketed table of gb_fast_kernel.
  union {
    struct {
      uint4 hbuckets[kFastBuckets];
      uint16_t hgid[kFastSlots32];
    };
    uint2 ktab[kBinSlots];
  };
  uint64_t bins[kBinCap * kFastGroups];
  uint64_t sum[kFastGroups];
  uint64_t mn[kFastGroups];
  uint64_t mx[kFastGroups];
  uint32_t tcnt[kFastGroups];
  uint32_t cnt[kFastGroups];
  uint32_t nulls[kFastGroups];
  uint32_t first[kFastGroups];
  uint32_t n_groups, need_insert, new_groups, deferred;
};
static_assert(kBinThreads == kFastGroups, "the reduction step gives every thread one group");
static_assert(sizeof(BinSmem) <= 227 * 1024, "BinSmem must fit the 227 KB a CTA can opt into");

// dense small keys (0..4095) map to themselves: no collisions at all for C3-like dictionaries; wider keys
// fold their upper bits in
__device__ __forceinline__ uint32_t bin_hash32(uint32_t k) { return (k ^ (k >> 12)) & uint32_t(kBinSlots - 1); }

__device__ __noinline__ int bin_lookup_far(const BinSmem& S, uint32_t k, uint32_t h) {
#pragma unroll 1
  for (int probe = 1; probe < 4 * kLocalProbe; ++probe) {
    h = (h + 1) & uint32_t(kBinSlots - 1);
    const uint2 e = S.ktab[h];
    if (e.x == k) return int(e.y);
    if (e.x == kEmpty32) return -1;
  }
  return -2;
}

// group id of a 32-bit key: -1 = not (yet) in the table, -2 = cannot be placed
__device__ __forceinline__ int bin_lookup32(const BinSmem& S, uint32_t k) {
  const uint32_t h = bin_hash32(k);
  const uint2 e = S.ktab[h];
  if (e.x == k) return int(e.y);   // the id, or kBinGidOverflow == -2
  if (e.x == kEmpty32) return -1;
  return bin_lookup_far(S, k, h);
}

// runs only between the two barriers of the insert phase; lookups resume after the second one
__device__ __forceinline__ void bin_insert32(BinSmem& S, uint32_t k) {
  uint32_t h = bin_hash32(k);
#pragma unroll 1
  for (int probe = 0; probe < 4 * kLocalProbe; ++probe) {
    uint32_t* kp = &S.ktab[h].x;
    uint32_t cur = *reinterpret_cast<volatile uint32_t*>(kp);
    if (cur == kEmpty32) {
      if (*reinterpret_cast<const volatile uint32_t*>(&S.n_groups) >= uint32_t(kFastGroups)) return;
      cur = atomicCAS(kp, kEmpty32, k);
      if (cur == kEmpty32) {
        const uint32_t g = atomicAdd(&S.n_groups, 1u);
        S.ktab[h].y = g < uint32_t(kFastGroups) ? g : kBinGidOverflow;
        return;
      }
    }
    if (cur == k) return;
    h = (h + 1) & uint32_t(kBinSlots - 1);
  }
}

template <int KMODE, bool VFAST>
__global__ void __launch_bounds__(kBinThreads, 1) gb_bin_kernel(const __grid_constant__ GbArgs a) {
  extern __shared__ __align__(16) unsigned char smem[];
  BinSmem& S = *reinterpret_cast<BinSmem*>(smem);
  const int t = threadIdx.x, lane = t & 31;
  constexpr bool KEY32 = KMODE == 1;
  // the raw columns of the NEXT tile are requested before this tile's barrier + reduction and decoded
  // after it: their latency hides behind work that needs few registers
  constexpr bool PREFETCH = KEY32 && VFAST;
  if (KEY32) {
    for (int i = t; i < kBinSlots; i += kBinThreads) S.ktab[i] = make_uint2(kEmpty32, 0xFFFFFFFFu);
  } else {
    for (int i = t; i < kFastBuckets; i += kBinThreads) S.hbuckets[i] = make_uint4(kEmpty32, kEmpty32, kEmpty32, kEmpty32);
    for (int i = t; i < kFastSlots32; i += kBinThreads) S.hgid[i] = uint16_t(kGidUnset);
  }
  {
    S.sum[t] = 0;
    S.mn[t] = ~uint64_t(0);
    S.mx[t] = 0;
    S.tcnt[t] = 0;
    S.cnt[t] = 0;
    S.nulls[t] = 0;
    S.first[t] = 0xFFFFFFFFu;
  }
  if (t == 0) {
    S.n_groups = 0;
    S.need_insert = 0;
    S.new_groups = 0;
    S.deferred = 0;
  }
  __syncthreads();
  // this launch serves the aggregates whose argument is column 0 (col == 0); a multi-column group-by runs
  // the kernel once per distinct argument column, the others carry col == -1
  int dom = kSigned;
  bool want_minmax = false;
  for (int j = a.s.n_aggs - 1; j >= 0; --j)
    if (a.s.aggs[j].col == 0) {
      dom = a.s.aggs[j].dom;
      if (a.s.aggs[j].kind == B200_HASH_MIN_MAX) want_minmax = true;
    }
  const int64_t begin = int64_t(blockIdx.x) * a.rows_per_block;
  const int64_t end = min(a.n, begin + a.rows_per_block);   // a.n is a multiple of kBinTile here
  uint64_t* zf = a.zf_scratch + size_t(blockIdx.x) * kFastGroups;

  // raw image of a tile's rows of this thread (PREFETCH only)
  uint4 rk[kFastSubs];
  ulonglong2 rva[kFastSubs], rvb[kFastSubs];
  uint32_t rvm[kFastSubs];
  auto request = [&](int64_t base) {
#pragma unroll
    for (int s = 0; s < kFastSubs; ++s) {
      const int64_t r0 = base + s * kBinSubRows + t * kFastSub;
      rk[s] = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint32_t*>(a.keys[0].data) + r0);
      const ulonglong2* vp = reinterpret_cast<const ulonglong2*>(reinterpret_cast<const uint64_t*>(a.cols[0].data) + r0);
      rva[s] = vp[0];
      rvb[s] = vp[1];
      rvm[s] = 0xFFu;
      if (a.cols[0].vv.words) {
        // r0 is a multiple of 4: its 4 validity bits never straddle more than two bytes
        const int64_t bit = a.cols[0].vv.bit_offset + r0;
        const uint8_t* bytes = reinterpret_cast<const uint8_t*>(a.cols[0].vv.words);
        rvm[s] = uint32_t(__ldg(reinterpret_cast<const uint16_t*>(bytes + ((bit >> 3) & ~int64_t(1))))) |
                 (uint32_t(__ldg(bytes + ((bit >> 3) & ~int64_t(1)) + 2)) << 16);
      }
    }
  };
  if (PREFETCH && begin < end) request(begin);

#pragma unroll 1
  for (int64_t base = begin; base < end; base += kBinTile) {
    uint64_t vbits[kBinRows];
    int gid[kBinRows];
    uint32_t vmask = 0;   // validity of the values
    uint32_t newkey = 0;  // rows whose key was not in the table when the tile started
    // ---- P1a: loads and group lookup
    {
      uint64_t key[kBinRows];
      uint32_t special = 0;
#pragma unroll
      for (int s = 0; s < kFastSubs; ++s) {
        const int64_t r0 = base + s * kBinSubRows + t * kFastSub;
        if (PREFETCH) {
          key[4 * s + 0] = rk[s].x; key[4 * s + 1] = rk[s].y; key[4 * s + 2] = rk[s].z; key[4 * s + 3] = rk[s].w;
          vbits[4 * s + 0] = rva[s].x; vbits[4 * s + 1] = rva[s].y; vbits[4 * s + 2] = rvb[s].x; vbits[4 * s + 3] = rvb[s].y;
          uint32_t m = 0xFu;
          if (a.cols[0].vv.words) {
            const int64_t bit = a.cols[0].vv.bit_offset + r0;
            m = rvm[s] >> (int(bit & 7) + 8 * int((bit >> 3) & 1));
          }
          vmask |= (m & 0xFu) << (4 * s);
#pragma unroll
          for (int i = 0; i < kFastSub; ++i)
            if (uint32_t(key[4 * s + i]) == kEmpty32) special |= 1u << (4 * s + i);
          continue;
        }
        if (KMODE == 1) {
          const uint4 k4 = *reinterpret_cast<const uint4*>(reinterpret_cast<const uint32_t*>(a.keys[0].data) + r0);
          key[4 * s + 0] = k4.x; key[4 * s + 1] = k4.y; key[4 * s + 2] = k4.z; key[4 * s + 3] = k4.w;
#pragma unroll
          for (int i = 0; i < kFastSub; ++i)
            if (uint32_t(key[4 * s + i]) == kEmpty32) special |= 1u << (4 * s + i);
        } else if (KMODE == 2) {
          const ulonglong2* kp = reinterpret_cast<const ulonglong2*>(reinterpret_cast<const uint64_t*>(a.keys[0].data) + r0);
          const ulonglong2 ka = kp[0], kb = kp[1];
          key[4 * s + 0] = ka.x; key[4 * s + 1] = ka.y; key[4 * s + 2] = kb.x; key[4 * s + 3] = kb.y;
#pragma unroll
          for (int i = 0; i < kFastSub; ++i)
            if (key[4 * s + i] == kEmpty) special |= 1u << (4 * s + i);
        } else {
#pragma unroll
          for (int i = 0; i < kFastSub; ++i) {
            int sp;
            key[4 * s + i] = load_key(a, r0 + i, &sp);
            if (sp) special |= 1u << (4 * s + i);
          }
        }
        if (VFAST) {
          const ulonglong2* vp = reinterpret_cast<const ulonglong2*>(reinterpret_cast<const uint64_t*>(a.cols[0].data) + r0);
          const ulonglong2 va = vp[0], vb = vp[1];
          vbits[4 * s + 0] = va.x; vbits[4 * s + 1] = va.y; vbits[4 * s + 2] = vb.x; vbits[4 * s + 3] = vb.y;
          uint32_t m = 0xFu;
          if (a.cols[0].vv.words) {
            const int64_t bit = a.cols[0].vv.bit_offset + r0;
            const uint8_t* bytes = reinterpret_cast<const uint8_t*>(a.cols[0].vv.words);
            const int sh = int(bit & 7);
            m = uint32_t(__ldg(bytes + (bit >> 3))) >> sh;
            if (sh > 4) m |= uint32_t(__ldg(bytes + (bit >> 3) + 1)) << (8 - sh);
          }
          vmask |= (m & 0xFu) << (4 * s);
        } else {
#pragma unroll
          for (int i = 0; i < kFastSub; ++i) {
            const Val v = load_val(a.cols[0], r0 + i);
            vbits[4 * s + i] = v.bits;
            vmask |= (v.valid ? 1u : 0u) << (4 * s + i);
          }
        }
      }
      uint32_t pending = 0;
#pragma unroll
      for (int i = 0; i < kBinRows; ++i) {
        if (a.dbg & 4) gid[i] = int(key[i] & 1023);
        else if (KEY32) gid[i] = bin_lookup32(S, uint32_t(key[i]));
        else gid[i] = fast_lookup<false>(S, key[i]);
        if (gid[i] == -1) pending |= 1u << i;
      }
      if (special) {   // the EMPTY-valued key cannot live in the table: its rows go to the generic kernel
#pragma unroll
        for (int i = 0; i < kBinRows; ++i)
          if ((special >> i) & 1u) gid[i] = -2;
        pending &= ~special;
      }
      newkey = pending;
      if (__any_sync(0xffffffffu, pending != 0) && lane == 0) S.need_insert = 1;
      __syncthreads();   // also orders the previous tile's reduction before this tile's bin writes
      if (S.need_insert) {   // CTA-uniform; taken for the first tile(s) only
#pragma unroll
        for (int i = 0; i < kBinRows; ++i)
          if ((pending >> i) & 1u) {
            if (KEY32) bin_insert32(S, uint32_t(key[i]));
            else fast_insert<false>(S, key[i]);
          }
        __syncthreads();
        if (t == 0) S.need_insert = 0;
#pragma unroll
        for (int i = 0; i < kBinRows; ++i)
          if ((pending >> i) & 1u) {
            gid[i] = KEY32 ? bin_lookup32(S, uint32_t(key[i])) : fast_lookup<false>(S, key[i]);
            if (gid[i] == -1) gid[i] = -2;
          }
        __syncwarp();
      }
    }
    // ---- P1b: slot in the group's bin (native shared atomic), straight into the bin
    uint32_t defer = 0, over = 0, zeros = 0;
#pragma unroll
    for (int i = 0; i < kBinRows; ++i) {
      const int g = gid[i];
      if (g < 0) {
        defer |= 1u << i;
        continue;
      }
      if (a.dbg & 2) continue;
      if ((vmask >> i) & 1u) {
        const uint32_t r = atomicAdd(&S.tcnt[g], 1u);
        if (r < uint32_t(kBinCap)) S.bins[r * kFastGroups + g] = vbits[i];
        else over |= 1u << i;
        if ((uint32_t(vbits[i] >> 32) & 0x7FFFFFFFu) == 0) zeros |= 1u << i;   // candidate +-0.0 (exact test below)
      } else {
        atomicAdd(&S.nulls[g], 1u);
      }
    }
    newkey &= ~defer;
    if (newkey) {   // first rows: only rows whose key entered the table in this tile can lower them
#pragma unroll
      for (int i = 0; i < kBinRows; ++i)
        if ((newkey >> i) & 1u) {
          const int64_t row = base + (i / kFastSub) * kBinSubRows + t * kFastSub + (i % kFastSub);
          atomicMin(&S.first[gid[i]], uint32_t(row - begin));
        }
    }
    if (zeros && want_minmax && dom == kFloat) {
      // +-0.0 ties in hash_min_max: the earliest zero row decides the sign (see accumulate())
#pragma unroll
      for (int i = 0; i < kBinRows; ++i)
        if (((zeros >> i) & 1u) && (vbits[i] << 1) == 0) {
          const int64_t row = base + (i / kFastSub) * kBinSubRows + t * kFastSub + (i % kFastSub);
          atomic_min_checked(zf + gid[i], ((a.row_base + uint64_t(row)) << 1) | (vbits[i] >> 63));
        }
    }
    if (defer) {
#pragma unroll
      for (int i = 0; i < kBinRows; ++i)
        if ((defer >> i) & 1u) {
          const int64_t row = base + (i / kFastSub) * kBinSubRows + t * kFastSub + (i % kFastSub);
          atomicOr(a.defer_out + (row >> 5), 1u << (row & 31));
        }
      atomicAdd(&S.deferred, uint32_t(__popc(defer)));
    }
    __syncwarp();
    if (__any_sync(0xffffffffu, over != 0)) {
      // ---- bin overflow (skewed keys): accumulate these rows directly.  Counts use the native 32-bit
      // atomic; float sums a CAS loop the whole warp stays converged in; extremes read-then-rarely-CAS.
#pragma unroll
      for (int i = 0; i < kBinRows; ++i) {
        const bool mine = (over >> i) & 1u;
        const int g = mine ? gid[i] : 0;
        if (mine) atomicAdd(&S.cnt[g], 1u);
        if (dom == kFloat) {
          bool pend = mine;
          while (__any_sync(0xffffffffu, pend)) {
            if (pend) {
              const uint64_t o = *reinterpret_cast<const volatile uint64_t*>(&S.sum[g]);
              const uint64_t nw = __double_as_longlong(__longlong_as_double(o) + __longlong_as_double(vbits[i]));
              if (atomicCAS(reinterpret_cast<unsigned long long*>(&S.sum[g]), (unsigned long long)o, (unsigned long long)nw) == o) pend = false;
            }
          }
        } else if (mine) {
          uint32_t* w = reinterpret_cast<uint32_t*>(&S.sum[g]);
          const uint32_t lo = uint32_t(vbits[i]), hi = uint32_t(vbits[i] >> 32);
          const uint32_t old_lo = atomicAdd(w, lo);
          const uint32_t carry = (old_lo + lo) < old_lo ? 1u : 0u;
          if (hi + carry) atomicAdd(w + 1, hi + carry);
        }
        if (want_minmax) {
          bool lt = false, gt = false;
          uint64_t o = 0;
          if (mine) {
            bool use = true;
            if (dom == kFloat) {
              const double dv = __longlong_as_double(vbits[i]);
              use = dv == dv;
              o = ord_f64(dv);
            } else {
              o = dom == kSigned ? (vbits[i] ^ 0x8000000000000000ull) : vbits[i];
            }
            if (use) {
              lt = o < *reinterpret_cast<const volatile uint64_t*>(&S.mn[g]);
              gt = o > *reinterpret_cast<const volatile uint64_t*>(&S.mx[g]);
            }
          }
          if (__any_sync(0xffffffffu, lt || gt)) {
            if (lt) atomicMin(reinterpret_cast<unsigned long long*>(&S.mn[g]), (unsigned long long)o);
            if (gt) atomicMax(reinterpret_cast<unsigned long long*>(&S.mx[g]), (unsigned long long)o);
            __syncwarp();
          }
        }
      }
    }
    if (PREFETCH && base + kBinTile < end) request(base + kBinTile);
    __syncthreads();
    // ---- P4: every thread reduces the bin of its own group with plain adds / compares
    {
      const int g = t;
      const uint32_t c_all = (a.dbg & 1) ? 0u : S.tcnt[g];
      if (c_all != 0) {
        const uint32_t c = min(c_all, uint32_t(kBinCap));
        S.tcnt[g] = 0;
        S.cnt[g] += c;
        const uint64_t* bin = S.bins + g;
        if (dom == kFloat) {
          // the group's running extremes ride along, so an update is the rare case and a value costs one
          // add and two compares (a NaN never compares true: it never wins, hash_aggregate.cc:672-673)
          const uint64_t omn = S.mn[g], omx = S.mx[g];
          double s = 0.0;
          double lo = omn == ~uint64_t(0) ? __longlong_as_double(0x7FF0000000000000ull) : unord_f64(omn);
          double hi = omx == 0 ? __longlong_as_double(0xFFF0000000000000ull) : unord_f64(omx);
#pragma unroll 4
          for (uint32_t j = 0; j < c; ++j) {
            const double v = __longlong_as_double(bin[j * kFastGroups]);
            s += v;
            if (v < lo || v > hi) {
              if (v < lo) lo = v;
              if (v > hi) hi = v;
            }
          }
          S.sum[g] = __double_as_longlong(__longlong_as_double(S.sum[g]) + s);
          const uint64_t olo = ord_f64(lo), ohi = ord_f64(hi);
          if (olo < omn) S.mn[g] = olo;
          if (ohi > omx) S.mx[g] = ohi;
        } else {
          const uint64_t flip = dom == kSigned ? 0x8000000000000000ull : 0ull;
          uint64_t s = 0, lo = ~uint64_t(0), hi = 0;
#pragma unroll 2
          for (uint32_t j = 0; j < c; ++j) {
            const uint64_t v = bin[j * kFastGroups];
            s += v;
            const uint64_t ov = v ^ flip;
            lo = ov < lo ? ov : lo;
            hi = ov > hi ? ov : hi;
          }
          S.sum[g] += s;
          if (lo < S.mn[g]) S.mn[g] = lo;
          if (hi > S.mx[g]) S.mx[g] = hi;
        }
      }
    }
    // no barrier here: the next tile's P1a touches neither tcnt nor the bins, and its own barrier
    // orders this reduction before the next tile's bin writes
  }
  __syncthreads();
  // ---- flush: every group of this CTA is folded into the global table
  for (int slot = t; slot < (KEY32 ? kBinSlots : kFastSlots64); slot += kBinThreads) {
    uint32_t g;
    uint64_t key;
    if (KEY32) {
      const uint2 e = S.ktab[slot];
      if (e.x == kEmpty32) continue;
      g = e.y;
      key = e.x;
    } else {
      g = S.hgid[slot];
      key = reinterpret_cast<const uint64_t*>(S.hbuckets)[slot];
    }
    if (g >= uint32_t(kFastGroups)) continue;   // unset or overflow
    bool claimed;
    uint64_t slot_index;
    char* gs = global_slot(a.t, a.s.stride, key, 0, true, &claimed, &slot_index);
    if (!gs) {
      gs = a.spill + size_t(atomicAdd(a.counters + 5, 1u)) * a.s.stride;
      init_slot(gs, a.s);
      reinterpret_cast<uint64_t*>(gs)[0] = key;
      reinterpret_cast<uint64_t*>(gs)[a.s.stride / 8 - 1] = 0;
      claimed = false;
    }
    if (claimed) atomicAdd(&S.new_groups, 1u);
    const uint64_t cnt = S.cnt[g], nulls = S.nulls[g];
    if (S.first[g] != 0xFFFFFFFFu) atomic_min_checked(reinterpret_cast<uint64_t*>(gs) + 1, a.row_base + uint64_t(begin) + S.first[g]);
    for (int j = 0; j < a.s.n_aggs; ++j) {
      const GbAgg& ag = a.s.aggs[j];
      if (ag.col != 0) continue;
      uint64_t* d = reinterpret_cast<uint64_t*>(gs + ag.off);
      if (ag.kind == B200_HASH_COUNT) {
        const uint64_t c = ag.option ? nulls : cnt;
        if (c) atomicAdd(reinterpret_cast<unsigned long long*>(d), (unsigned long long)c);
      } else if (ag.kind == B200_HASH_MIN_MAX) {
        if (cnt) {
          atomic_min_checked(d + 0, S.mn[g]);
          atomic_max_checked(d + 1, S.mx[g]);
          atomic_min_checked(d + 2, zf[g]);
        }
        const uint64_t flags = (cnt ? 1ull : 0ull) | (nulls ? 2ull : 0ull);
        if (flags) atomic_or_checked(d + 3, flags);
      } else if (cnt) {
        if (ag.dom == kFloat) atomicAdd(reinterpret_cast<double*>(d), __longlong_as_double(S.sum[g]));
        else atomicAdd(reinterpret_cast<unsigned long long*>(d), (unsigned long long)S.sum[g]);
        atomicAdd(reinterpret_cast<unsigned long long*>(d + 1), (unsigned long long)cnt);
      }
    }
  }
  __syncthreads();
  if (t == 0) {
    if (S.new_groups) atomicAdd(a.counters + 0, S.new_groups);
    if (S.deferred) atomicAdd(a.counters + 1, S.deferred);
  }
}

// ---- binned path, pipelined form for the C3 shape (one <= 32-bit key column without nulls, one 64-bit
// value column).  gb_bin_kernel with no work at all in its row loop still needed 4.1 ms per 1 B rows
// (2.9 TB/s): with one bulk-synchronous CTA per SM the column loads of a tile were requested only after
// the previous tile's loads had been consumed.  Here the loads run TWO half-tiles (2 x 4096 rows, 2 x 12
// registers per thread) ahead of the work and flow across the barriers, and the per-tile consensus barrier
// for new keys is gone: a row whose key is not in the table yet inserts it on the spot (CAS on the key
// word, then the dense id is published in the same 8-byte slot; concurrent lookups of that key spin for
// the few cycles until it is).  Every slot remembers the tile it was born in, because a row that HITS a
// key born in the current tile may still be the group's first row (warps are not ordered inside a tile).
constexpr uint32_t kB32Unset = 0xFFFFu, kB32Overflow = 0xFFFEu;
// whole-word forms of the two markers: a published id word is (birth tile << 16) | id with id < 1024, so
// "found, placed and born before tile T" is the single unsigned compare  word < (T << 16)  (tiles only grow)
constexpr uint32_t kB32UnsetWord = 0xFFFFFFFFu, kB32OverflowWord = 0xFFFFFFFEu;

struct Raw32 {
  uint4 k;
  ulonglong2 va, vb;
  uint16_t vm0;   // the 16-bit validity word holding the four rows' bits (bit offset % 4 == 0, see the dispatch)
};

// key -> dense group id (>= 0), or -2 when the CTA's table is full.  *fresh = the group was born in `tile`.
__device__ __forceinline__ int b32_find_or_insert(BinSmem& S, uint32_t k, uint32_t tile, bool* fresh) {
  uint32_t h = bin_hash32(k);
  uint2 e = S.ktab[h];
  if (e.x != k || (e.y & 0xFFFFu) >= kB32Overflow) {
    // slow path: probe, insert, or wait for the inserting thread to publish the id
#pragma unroll 1
    for (int probe = 0; probe < 4 * kLocalProbe; ++probe) {
      uint32_t cur = *reinterpret_cast<volatile uint32_t*>(&S.ktab[h].x);
      if (cur == kEmpty32) {
        if (*reinterpret_cast<const volatile uint32_t*>(&S.n_groups) >= uint32_t(kFastGroups)) return -2;
        cur = atomicCAS(&S.ktab[h].x, kEmpty32, k);
        if (cur == kEmpty32) {
          const uint32_t g = atomicAdd(&S.n_groups, 1u);
          const uint32_t y = g < uint32_t(kFastGroups) ? (g | (tile << 16)) : kB32OverflowWord;
          *reinterpret_cast<volatile uint32_t*>(&S.ktab[h].y) = y;
          *fresh = true;
          return g < uint32_t(kFastGroups) ? int(g) : -2;
        }
      }
      if (cur == k) {
        uint32_t y;
        do {
          y = *reinterpret_cast<volatile uint32_t*>(&S.ktab[h].y);
        } while ((y & 0xFFFFu) == kB32Unset);
        e.y = y;
        break;
      }
      h = (h + 1) & uint32_t(kBinSlots - 1);
      e.y = kB32Overflow;   // probe limit reached without finding the key: cannot be placed
    }
  }
  const uint32_t g = e.y & 0xFFFFu;
  *fresh = (e.y >> 16) == tile;
  return g < uint32_t(kFastGroups) ? int(g) : -2;
}

__global__ void __launch_bounds__(kBinThreads, 1) gb_bin32_kernel(const __grid_constant__ GbArgs a) {
  extern __shared__ __align__(16) unsigned char smem[];
  BinSmem& S = *reinterpret_cast<BinSmem*>(smem);
  const int t = threadIdx.x;
  for (int i = t; i < kBinSlots; i += kBinThreads) S.ktab[i] = make_uint2(kEmpty32, kB32UnsetWord);
  S.sum[t] = 0;
  S.mn[t] = ~uint64_t(0);
  S.mx[t] = 0;
  S.tcnt[t] = 0;
  S.cnt[t] = 0;
  S.nulls[t] = 0;
  S.first[t] = 0xFFFFFFFFu;
  if (t == 0) {
    S.n_groups = 0;
    S.new_groups = 0;
    S.deferred = 0;
  }
  __syncthreads();
  // this launch serves the aggregates whose argument is column 0 (col == 0); a multi-column group-by runs
  // the kernel once per distinct argument column, the others carry col == -1
  int dom = kSigned;
  bool want_minmax = false;
  for (int j = a.s.n_aggs - 1; j >= 0; --j)
    if (a.s.aggs[j].col == 0) {
      dom = a.s.aggs[j].dom;
      if (a.s.aggs[j].kind == B200_HASH_MIN_MAX) want_minmax = true;
    }
  const int64_t begin = int64_t(blockIdx.x) * a.rows_per_block;
  const int64_t end = min(a.n, begin + a.rows_per_block);   // a.n is a multiple of kBinTile here
  uint64_t* zf = a.zf_scratch + size_t(blockIdx.x) * kFastGroups;
  const uint32_t* kcol = reinterpret_cast<const uint32_t*>(a.keys[0].data);
  const uint64_t* vcol = reinterpret_cast<const uint64_t*>(a.cols[0].data);
  const uint8_t* vbytes = reinterpret_cast<const uint8_t*>(a.cols[0].vv.words);
  const int64_t voff = a.cols[0].vv.bit_offset;

  // half-tile h (4096 rows) of this CTA starts at row begin + h * kBinSubRows; this thread owns 4 rows of it
  // (no branch around the loads: past the CTA's last row the thread re-reads its rows of the first half-tile and
  // nobody looks at them.  With `if (r0 < end)` the two paths met in registers of the compiler's choosing and the
  // loaded values were MOVED there right behind the loads: the profile showed the whole memory latency on those
  // two moves, 20 % of the kernel's samples)
  auto request = [&](Raw32& r, int64_t r0) {
    r0 = r0 < end ? r0 : begin + int64_t(t) * kFastSub;
    {
#if B200_BIN_LD == 0
      r.k = *reinterpret_cast<const uint4*>(kcol + r0);
      const ulonglong2* vp = reinterpret_cast<const ulonglong2*>(vcol + r0);
      r.va = vp[0];
      r.vb = vp[1];
#else
      asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                   : "=r"(r.k.x), "=r"(r.k.y), "=r"(r.k.z), "=r"(r.k.w) : "l"(kcol + r0));
#if B200_BIN_LD == 2
      asm volatile("ld.global.nc.L1::no_allocate.v4.u64 {%0,%1,%2,%3}, [%4];"
                   : "=l"(r.va.x), "=l"(r.va.y), "=l"(r.vb.x), "=l"(r.vb.y) : "l"(vcol + r0));
#else
      asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0,%1}, [%2];" : "=l"(r.va.x), "=l"(r.va.y) : "l"(vcol + r0));
      asm volatile("ld.global.nc.L1::no_allocate.v2.u64 {%0,%1}, [%2];" : "=l"(r.vb.x), "=l"(r.vb.y) : "l"(vcol + r0 + 2));
#endif
#endif
      if (vbytes) {
        // r0 and the bit offset are multiples of 4: the 4 validity bits lie inside one 16-bit aligned word
        const int64_t byte = ((voff + r0) >> 3) & ~int64_t(1);
        r.vm0 = __ldg(reinterpret_cast<const uint16_t*>(vbytes + byte));
      }
    }
  };
  // one half-tile of this thread's rows: lookup / insert, slot in the bin, rare side paths
  auto work = [&](const Raw32& r, int64_t r0, uint32_t tile) {
    if (a.dbg & 8) {   // timing experiment: loads only
      if ((r.k.x ^ r.k.y ^ r.k.z ^ r.k.w ^ uint32_t(r.va.x ^ r.va.y ^ r.vb.x ^ r.vb.y) ^ r.vm0) == 0x12345678u) S.deferred = 1;
      return;
    }
    const uint32_t key[4] = {r.k.x, r.k.y, r.k.z, r.k.w};
    const uint64_t vbits[4] = {r.va.x, r.va.y, r.vb.x, r.vb.y};
    uint32_t vmask = 0xFu;
    if (vbytes) {
      const int64_t bit = voff + r0;
      vmask = (uint32_t(r.vm0) >> int(bit & 15)) & 0xFu;
    }
    int gid[4];
    uint32_t fresh = 0, defer = 0, over = 0, zeros = 0;
    // steady state (every tile but the first few): the four keys are in the table, born in an earlier tile, and
    // the four values are valid.  Four independent probes, two compares per row, then four independent rank
    // atomics: no branch between them, nothing of the insert / first-row / null machinery below is touched
    // (the generic form spent ~50 instructions per row here, half of them branch and re-convergence overhead)
    bool steady;
    {
      uint2 e[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) e[i] = S.ktab[bin_hash32(key[i])];
      const uint32_t born_before = tile << 16;
      uint32_t miss = 0;   // no short-circuit: the four loads are in flight together
#pragma unroll
      for (int i = 0; i < 4; ++i) miss |= (e[i].x ^ key[i]) | uint32_t(e[i].y >= born_before);
      steady = miss == 0;
      if (steady) {
        // a null row counts in nulls[] instead of taking a rank: same instruction, other counter array, so that
        // lanes with and without nulls stay on one path (1 % nulls puts a null into 3 of 4 warps)
        uint32_t rk[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          gid[i] = int(e[i].y & 0xFFFFu);
          const bool valid = (vmask >> i) & 1u;
          const uint32_t r = atomicAdd(valid ? &S.tcnt[gid[i]] : &S.nulls[gid[i]], 1u);
          rk[i] = valid ? r : 0u;
        }
        uint32_t mag = 0xFFFFFFFFu;   // smallest |upper word| of the valid values: 0 <=> one of them may be +-0.0
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          if ((vmask >> i) & 1u) {
            if (rk[i] < uint32_t(kBinCap)) S.bins[rk[i] * kFastGroups + gid[i]] = vbits[i];
            mag = min(mag, uint32_t(vbits[i] >> 32) & 0x7FFFFFFFu);
          }
        }
        if ((max(max(rk[0], rk[1]), max(rk[2], rk[3])) >= uint32_t(kBinCap)) | (mag == 0)) {   // rare
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            if (rk[i] >= uint32_t(kBinCap)) over |= 1u << i;
            if (((vmask >> i) & 1u) && (uint32_t(vbits[i] >> 32) & 0x7FFFFFFFu) == 0) zeros |= 1u << i;   // candidate +-0.0
          }
        }
      }
    }
    if (!steady) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        bool f = false;
        gid[i] = key[i] == kEmpty32 ? -2 : b32_find_or_insert(S, key[i], tile, &f);
        if (f) fresh |= 1u << i;
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int g = gid[i];
        if (g < 0) {
          defer |= 1u << i;
          continue;
        }
        if ((vmask >> i) & 1u) {
          const uint32_t rk = atomicAdd(&S.tcnt[g], 1u);
          if (rk < uint32_t(kBinCap)) S.bins[rk * kFastGroups + g] = vbits[i];
          else over |= 1u << i;
          if ((uint32_t(vbits[i] >> 32) & 0x7FFFFFFFu) == 0) zeros |= 1u << i;   // candidate +-0.0
        } else {
          atomicAdd(&S.nulls[g], 1u);
        }
      }
    }
    fresh &= ~defer;
    if (fresh) {
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if ((fresh >> i) & 1u) atomicMin(&S.first[gid[i]], uint32_t(r0 + i - begin));
    }
    if (zeros && want_minmax && dom == kFloat) {
      // +-0.0 ties in hash_min_max: the earliest zero row decides the sign (see accumulate())
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (((zeros >> i) & 1u) && (vbits[i] << 1) == 0)
          atomic_min_checked(zf + gid[i], ((a.row_base + uint64_t(r0 + i)) << 1) | (vbits[i] >> 63));
    }
    if (defer) {
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if ((defer >> i) & 1u) atomicOr(a.defer_out + ((r0 + i) >> 5), 1u << ((r0 + i) & 31));
      atomicAdd(&S.deferred, uint32_t(__popc(defer)));
    }
    if (over) {
      // bin overflow (skewed keys): accumulate directly.  Counts: native 32-bit atomic; float sums: CAS loop
      // (lanes may retry independently here: this path is already the slow one); extremes: checked atomics
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (!((over >> i) & 1u)) continue;
        const int g = gid[i];
        atomicAdd(&S.cnt[g], 1u);
        if (dom == kFloat) {
          unsigned long long* sp = reinterpret_cast<unsigned long long*>(&S.sum[g]);
          unsigned long long o = *reinterpret_cast<volatile unsigned long long*>(sp), seen;
          do {
            seen = o;
            o = atomicCAS(sp, seen, (unsigned long long)__double_as_longlong(__longlong_as_double(seen) + __longlong_as_double(vbits[i])));
          } while (o != seen);
        } else {
          uint32_t* w = reinterpret_cast<uint32_t*>(&S.sum[g]);
          const uint32_t lo = uint32_t(vbits[i]), hi = uint32_t(vbits[i] >> 32);
          const uint32_t old_lo = atomicAdd(w, lo);
          const uint32_t carry = (old_lo + lo) < old_lo ? 1u : 0u;
          if (hi + carry) atomicAdd(w + 1, hi + carry);
        }
        if (want_minmax) {
          uint64_t o;
          bool use = true;
          if (dom == kFloat) {
            const double dv = __longlong_as_double(vbits[i]);
            use = dv == dv;
            o = ord_f64(dv);
          } else {
            o = dom == kSigned ? (vbits[i] ^ 0x8000000000000000ull) : vbits[i];
          }
          if (use) {
            atomic_min_checked(&S.mn[g], o);
            atomic_max_checked(&S.mx[g], o);
          }
        }
      }
    }
  };

  // The demand loads above are issued at most one tile ahead (registers), and the profile of that form shows the
  // work phase waiting for them (long scoreboard on the first use of a half-tile, DRAM at 28 %).  Six threads
  // therefore ask the copy engine to pull the tile after next (.. B200_BIN_PF tiles ahead) into L2, 16 KB each: no
  // registers, no shared memory, no entry in the L1 miss queue; the demand load then only has L2 latency to cover.
  auto prefetch_tile = [&](int64_t base) {
    if (B200_BIN_PF > 0 && t < 6 && base < end) {
      const void* p = t < 2 ? static_cast<const void*>(kcol + base + t * 4096) : static_cast<const void*>(vcol + base + (t - 2) * 2048);
      asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(16384) : "memory");
    }
  };
  for (int d = 1; d <= B200_BIN_PF; ++d) prefetch_tile(begin + int64_t(d) * kBinTile);
  // ONE load site per half-tile buffer: the loop starts one tile early with the work switched off, so that the
  // first requests are issued by the same instructions as all later ones (a separate prologue request landed in
  // other registers and cost a move of freshly loaded values, i.e. the whole memory latency, in every iteration)
  Raw32 ra, rb;
  uint32_t tile = 0;
#pragma unroll 1
  for (int64_t base = begin - kBinTile; base < end; base += kBinTile) {
    const bool live = base >= begin;   // uniform
    const int64_t r0 = base + int64_t(t) * kFastSub;
    prefetch_tile(base + int64_t(1 + B200_BIN_PF) * kBinTile);
    // the next request goes into the registers the work has just consumed (a copy of the half-tile kept
    // beside two outstanding requests did not fit 64 registers: 48 B of stack per thread, i.e. spills to L2
    // with 205 KB of the SM's memory taken by the bins).  Its lead is still the other half-tile's work, the
    // reduction and both barriers
    if (live) work(ra, r0, tile & 0xFFFFu);
    request(ra, r0 + 2 * kBinSubRows);
    if (live) work(rb, r0 + kBinSubRows, tile & 0xFFFFu);
    request(rb, r0 + 3 * kBinSubRows);
    if (!live) continue;
    ++tile;
    __syncthreads();
    // ---- every thread reduces the bin of its own group with plain adds / compares
    {
      const int g = t;
      const uint32_t c_all = S.tcnt[g];
      if (c_all != 0) {
        const uint32_t c = min(c_all, uint32_t(kBinCap));
        S.tcnt[g] = 0;
        S.cnt[g] += c;
        const uint64_t* bin = S.bins + g;
        if (dom == kFloat) {
          const uint64_t omn = S.mn[g], omx = S.mx[g];
          double s = 0.0;
          double lo = omn == ~uint64_t(0) ? __longlong_as_double(0x7FF0000000000000ull) : unord_f64(omn);
          double hi = omx == 0 ? __longlong_as_double(0xFFF0000000000000ull) : unord_f64(omx);
#pragma unroll 4
          for (uint32_t j = 0; j < c; ++j) {
            const double v = __longlong_as_double(bin[j * kFastGroups]);
            s += v;
            if (v < lo || v > hi) {   // a NaN never compares true: it never wins (hash_aggregate.cc:672-673)
              if (v < lo) lo = v;
              if (v > hi) hi = v;
            }
          }
          S.sum[g] = __double_as_longlong(__longlong_as_double(S.sum[g]) + s);
          const uint64_t olo = ord_f64(lo), ohi = ord_f64(hi);
          if (olo < omn) S.mn[g] = olo;
          if (ohi > omx) S.mx[g] = ohi;
        } else {
          const uint64_t flip = dom == kSigned ? 0x8000000000000000ull : 0ull;
          uint64_t s = 0, lo = ~uint64_t(0), hi = 0;
#pragma unroll 2
          for (uint32_t j = 0; j < c; ++j) {
            const uint64_t v = bin[j * kFastGroups];
            s += v;
            const uint64_t ov = v ^ flip;
            lo = ov < lo ? ov : lo;
            hi = ov > hi ? ov : hi;
          }
          S.sum[g] += s;
          if (lo < S.mn[g]) S.mn[g] = lo;
          if (hi > S.mx[g]) S.mx[g] = hi;
        }
      }
    }
    __syncthreads();   // the bins and their counters are free again
  }
  // ---- flush: every group of this CTA is folded into the global table
  for (int slot = t; slot < kBinSlots; slot += kBinThreads) {
    const uint2 e = S.ktab[slot];
    if (e.x == kEmpty32) continue;
    const uint32_t g = e.y & 0xFFFFu;
    if (g >= uint32_t(kFastGroups)) continue;   // overflow marker
    const uint64_t key = e.x;
    bool claimed;
    uint64_t slot_index;
    char* gs = global_slot(a.t, a.s.stride, key, 0, true, &claimed, &slot_index);
    if (!gs) {
      gs = a.spill + size_t(atomicAdd(a.counters + 5, 1u)) * a.s.stride;
      init_slot(gs, a.s);
      reinterpret_cast<uint64_t*>(gs)[0] = key;
      reinterpret_cast<uint64_t*>(gs)[a.s.stride / 8 - 1] = 0;
      claimed = false;
    }
    if (claimed) atomicAdd(&S.new_groups, 1u);
    const uint64_t cnt = S.cnt[g], nulls = S.nulls[g];
    if (S.first[g] != 0xFFFFFFFFu) atomic_min_checked(reinterpret_cast<uint64_t*>(gs) + 1, a.row_base + uint64_t(begin) + S.first[g]);
    for (int j = 0; j < a.s.n_aggs; ++j) {
      const GbAgg& ag = a.s.aggs[j];
      if (ag.col != 0) continue;
      uint64_t* d = reinterpret_cast<uint64_t*>(gs + ag.off);
      if (ag.kind == B200_HASH_COUNT) {
        const uint64_t c = ag.option ? nulls : cnt;
        if (c) atomicAdd(reinterpret_cast<unsigned long long*>(d), (unsigned long long)c);
      } else if (ag.kind == B200_HASH_MIN_MAX) {
        if (cnt) {
          atomic_min_checked(d + 0, S.mn[g]);
          atomic_max_checked(d + 1, S.mx[g]);
          atomic_min_checked(d + 2, zf[g]);
        }
        const uint64_t flags = (cnt ? 1ull : 0ull) | (nulls ? 2ull : 0ull);
        if (flags) atomic_or_checked(d + 3, flags);
      } else if (cnt) {
        if (ag.dom == kFloat) atomicAdd(reinterpret_cast<double*>(d), __longlong_as_double(S.sum[g]));
        else atomicAdd(reinterpret_cast<unsigned long long*>(d), (unsigned long long)S.sum[g]);
        atomicAdd(reinterpret_cast<unsigned long long*>(d + 1), (unsigned long long)cnt);
      }
    }
  }
  __syncthreads();
  if (t == 0) {
    if (S.new_groups) atomicAdd(a.counters + 0, S.new_groups);
    if (S.deferred) atomicAdd(a.counters + 1, S.deferred);
  }
}

// ---- tiny-cardinality path (<= 8 groups: the Q1 shape) ------------------------------------------
// With a handful of groups every shared-memory counter or bin is hammered by the whole CTA (measured on
// Q1's 4 groups: 24 ms per column with the counting-sort kernel, 180 ms with the binned one).  Here nothing
// is shared while rows stream: every thread owns a private (sum, count) accumulator per group in shared
// memory, laid out [group][thread] so a warp's accesses are conflict-free, and updates them with plain
// loads / adds / stores.  Keys are matched against a <= 8-entry CTA table (linear scan, all lanes read the
// same word: a broadcast; the table is re-read into registers once per 4 rows), new keys are claimed with one CAS.  Rare events (null value, first row of a
// group in a thread) use shared atomics.  At the end the thread-private partials are tree-reduced per
// group and folded into the global table.  hash_sum / hash_mean / hash_count only (no min_max state).
constexpr int kTinyGroups = 8;
constexpr int kTinyThreads = 256;

struct TinySmem {
  uint64_t sum[kTinyGroups][kTinyThreads];
  uint32_t cnt[kTinyGroups][kTinyThreads];
  uint64_t keys[kTinyGroups];
  uint32_t nulls[kTinyGroups];
  uint32_t first[kTinyGroups];
  uint64_t red_sum[kTinyThreads / 32];
  uint32_t red_cnt[kTinyThreads / 32];
  uint32_t new_groups, deferred;
};

// the CTA's thread-private partials S.sum / S.cnt [group][thread] -> tree reduction per group -> one thread folds the
// group into the global table (shared by gb_tiny_kernel and gb_tiny8_kernel; the caller has synchronised)
__device__ __forceinline__ void tiny_fold(const GbArgs& a, TinySmem& S, const int dom, const int64_t begin) {
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  // ---- per group: tree reduction of the thread-private partials, then one thread folds it into the table
  for (int g = 0; g < kTinyGroups; ++g) {
    const uint64_t key = S.keys[g];
    if (key == kEmpty) break;   // entries fill in order
    if (S.first[g] == 0xFFFFFFFFu) continue;   // CTA-uniform: no (selected) row of this CTA belongs to the key
    uint64_t sv = S.sum[g][t];
    uint32_t cv = S.cnt[g][t];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
      const uint64_t os = __shfl_down_sync(0xffffffffu, sv, d);
      const uint32_t oc = __shfl_down_sync(0xffffffffu, cv, d);
      if (dom == kFloat) sv = __double_as_longlong(__longlong_as_double(sv) + __longlong_as_double(os));
      else sv += os;
      cv += oc;
    }
    if (lane == 0) {
      S.red_sum[warp] = sv;
      S.red_cnt[warp] = cv;
    }
    __syncthreads();
    if (t == 0) {
      uint64_t ts = S.red_sum[0];
      uint64_t tc = S.red_cnt[0];
      for (int w = 1; w < kTinyThreads / 32; ++w) {
        if (dom == kFloat) ts = __double_as_longlong(__longlong_as_double(ts) + __longlong_as_double(S.red_sum[w]));
        else ts += S.red_sum[w];
        tc += S.red_cnt[w];
      }
      bool claimed;
      uint64_t slot_index;
      char* gs = global_slot(a.t, a.s.stride, key, 0, true, &claimed, &slot_index);
      if (!gs) {
        gs = a.spill + size_t(atomicAdd(a.counters + 5, 1u)) * a.s.stride;
        init_slot(gs, a.s);
        reinterpret_cast<uint64_t*>(gs)[0] = key;
        reinterpret_cast<uint64_t*>(gs)[a.s.stride / 8 - 1] = 0;
        claimed = false;
      }
      if (claimed) S.new_groups += 1;
      const uint64_t nulls = S.nulls[g];
      if (S.first[g] != 0xFFFFFFFFu) atomic_min_checked(reinterpret_cast<uint64_t*>(gs) + 1, a.row_base + uint64_t(begin) + S.first[g]);
      for (int j = 0; j < a.s.n_aggs; ++j) {
        const GbAgg& ag = a.s.aggs[j];
        if (ag.col != 0) continue;
        uint64_t* d = reinterpret_cast<uint64_t*>(gs + ag.off);
        if (ag.kind == B200_HASH_COUNT) {
          const uint64_t c = ag.option ? nulls : tc;
          if (c) atomicAdd(reinterpret_cast<unsigned long long*>(d), (unsigned long long)c);
        } else if (tc) {   // hash_sum / hash_mean (this path is not chosen for hash_min_max)
          if (ag.dom == kFloat) atomicAdd(reinterpret_cast<double*>(d), __longlong_as_double(ts));
          else atomicAdd(reinterpret_cast<unsigned long long*>(d), (unsigned long long)ts);
          atomicAdd(reinterpret_cast<unsigned long long*>(d + 1), (unsigned long long)tc);
        }
      }
    }
    __syncthreads();
  }
  if (t == 0) {
    if (S.new_groups) atomicAdd(a.counters + 0, S.new_groups);
    if (S.deferred) atomicAdd(a.counters + 1, S.deferred);
  }
}

__global__ void __launch_bounds__(kTinyThreads, 4) gb_tiny_kernel(const __grid_constant__ GbArgs a) {
  extern __shared__ __align__(16) unsigned char smem[];
  TinySmem& S = *reinterpret_cast<TinySmem*>(smem);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
#pragma unroll
  for (int g = 0; g < kTinyGroups; ++g) {
    S.sum[g][t] = 0;
    S.cnt[g][t] = 0;
  }
  if (t < kTinyGroups) {
    S.keys[t] = kEmpty;
    S.nulls[t] = 0;
    S.first[t] = 0xFFFFFFFFu;
  }
  if (t == 0) {
    S.new_groups = 0;
    S.deferred = 0;
  }
  __syncthreads();
  int dom = kSigned;
  for (int j = a.s.n_aggs - 1; j >= 0; --j)
    if (a.s.aggs[j].col == 0) dom = a.s.aggs[j].dom;
  const int64_t begin = int64_t(blockIdx.x) * a.rows_per_block;
  const int64_t end = min(a.n, begin + a.rows_per_block);
  uint32_t seen = 0;   // groups this thread has met (their first row in this thread is already recorded)
  constexpr int kTinyUnroll = 4;   // rows in flight per thread: all loads of an iteration are issued first
#pragma unroll 1
  for (int64_t row0 = begin + t; row0 < end; row0 += kTinyUnroll * kTinyThreads) {
    uint64_t keyv[kTinyUnroll];
    Val val[kTinyUnroll];
    int spec[kTinyUnroll];
#pragma unroll
    for (int u = 0; u < kTinyUnroll; ++u) {
      const int64_t row = row0 + u * kTinyThreads;
      spec[u] = 1;
      keyv[u] = 0;
      val[u].bits = 0;
      val[u].valid = false;
      if (row < end) {
        keyv[u] = load_key(a, row, &spec[u]);
        val[u] = load_val(a.cols[0], row);
      }
    }
    uint64_t tk[kTinyGroups];   // the CTA's key table, as of now (entries only ever go from EMPTY to a key)
#pragma unroll
    for (int q = 0; q < kTinyGroups; ++q) tk[q] = *reinterpret_cast<const volatile uint64_t*>(&S.keys[q]);
#pragma unroll
    for (int u = 0; u < kTinyUnroll; ++u) {
      const int64_t row = row0 + u * kTinyThreads;
      if (row >= end) break;
      if (a.select && !((a.select[row >> 5] >> (row & 31)) & 1u)) continue;
      const uint64_t key = keyv[u];
      int g = -1;
      if (!spec[u]) {
#pragma unroll
        for (int q = 0; q < kTinyGroups; ++q)
          if (tk[q] == key) g = q;
        if (g < 0) {   // not in the snapshot: claim the first free entry, or find the key claimed meanwhile
#pragma unroll 1
          for (int q = 0; q < kTinyGroups && g < 0; ++q) {
            const unsigned long long old = atomicCAS(reinterpret_cast<unsigned long long*>(&S.keys[q]), (unsigned long long)kEmpty, (unsigned long long)key);
            if (old == kEmpty || old == key) g = q;
          }
#pragma unroll
          for (int q = 0; q < kTinyGroups; ++q) tk[q] = *reinterpret_cast<const volatile uint64_t*>(&S.keys[q]);
        }
      }
      if (g < 0) {   // special key or more than 8 groups in this CTA: the generic kernel takes the row
        atomicOr(a.defer_out + (row >> 5), 1u << (row & 31));
        atomicAdd(&S.deferred, 1u);
        continue;
      }
      if (!((seen >> g) & 1u)) {
        seen |= 1u << g;
        atomicMin(&S.first[g], uint32_t(row - begin));   // rows of a thread come in increasing order
      }
      if (val[u].valid) {
        if (dom == kFloat) S.sum[g][t] = __double_as_longlong(__longlong_as_double(S.sum[g][t]) + __longlong_as_double(val[u].bits));
        else S.sum[g][t] += val[u].bits;
        S.cnt[g][t] += 1;
      } else {
        atomicAdd(&S.nulls[g], 1u);
      }
    }
  }
  __syncthreads();
  tiny_fold(a, S, dom, begin);
}

// ---- <= 8 groups, byte-wide keys, one 64-bit value column: register accumulators ---------------------
// The Q1 shape (two uint8 keys, float64 measures).  gb_tiny_kernel's generic loaders (a type switch and a bitmap
// probe per key column and per value, 64-bit key compares) cost ~250 instructions per row; here the packed key
// is at most 24 bits, the CTA's key table lives in 8 registers, and every thread keeps the 8 groups' sums and
// counts in REGISTERS, updated with predicated adds (no shared memory while rows stream; a predicated add, not an
// add of zero, so a group of -0.0 values keeps its sign).  The epilogue is gb_tiny_kernel's: the accumulators are
// parked in its [group][thread] shared arrays, tree-reduced and folded into the global table.
template <bool kFloatSum, bool kSelect>
__global__ void __launch_bounds__(kTinyThreads, 4) gb_tiny8_kernel(const __grid_constant__ GbArgs a) {
  extern __shared__ __align__(16) unsigned char smem[];
  TinySmem& S = *reinterpret_cast<TinySmem*>(smem);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  if (t < kTinyGroups) {
    S.keys[t] = kEmpty;
    S.nulls[t] = 0;
    S.first[t] = 0xFFFFFFFFu;
  }
  if (t == 0) {
    S.new_groups = 0;
    S.deferred = 0;
  }
  __syncthreads();
  const int64_t begin = int64_t(blockIdx.x) * a.rows_per_block;
  const int64_t end = min(a.n, begin + a.rows_per_block);
  const uint8_t* k0 = reinterpret_cast<const uint8_t*>(a.keys[0].data);
  const uint8_t* k1 = a.s.n_keys > 1 ? reinterpret_cast<const uint8_t*>(a.keys[1].data) : nullptr;
  const uint8_t* k2 = a.s.n_keys > 2 ? reinterpret_cast<const uint8_t*>(a.keys[2].data) : nullptr;
  const uint64_t* vals = reinterpret_cast<const uint64_t*>(a.cols[0].data);
  const BitmapRef vv = a.cols[0].vv;
  uint32_t tk[kTinyGroups];      // low words of the CTA's key table (EMPTY reads as 0xFFFFFFFF, above any 24-bit key)
  uint64_t acc[kTinyGroups];
  uint32_t cnt[kTinyGroups];
#pragma unroll
  for (int q = 0; q < kTinyGroups; ++q) {
    tk[q] = 0xFFFFFFFFu;
    acc[q] = 0;
    cnt[q] = 0;
  }
  uint32_t seen = 0;
  constexpr int kU = 4;   // rows in flight per thread
#pragma unroll 1
  for (int64_t row0 = begin + t; row0 < end; row0 += kU * kTinyThreads) {
    uint32_t key[kU];
    uint64_t v[kU];
    bool valid[kU];
    uint32_t sel[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t row = min(row0 + u * kTinyThreads, end - 1);   // clamped: the loads need no branch
      sel[u] = kSelect ? __ldg(a.select + (row >> 5)) : 0xffffffffu;
      uint32_t k = k0[row];
      if (k1) k |= uint32_t(k1[row]) << 8;
      if (k2) k |= uint32_t(k2[row]) << 16;
      key[u] = k;
      v[u] = vals[row];
      valid[u] = bitmap_get(vv, row);
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t row = row0 + u * kTinyThreads;
      // A row the selection leaves out takes the same path as every other row (its key is matched, even entered
      // into the CTA's table) and only its side effects are predicated off: a warp in which a few lanes skipped the
      // row body ran 8x slower (measured with a 96 % selection; 0 %, 50 % and 100 % were fine).  A key that only
      // unselected rows brought in has no first row and is dropped by tiny_fold.
      if (row < end) {
        const bool on = !kSelect || ((sel[u] >> (row & 31)) & 1u);
        int g = -1;
#pragma unroll
        for (int q = 0; q < kTinyGroups; ++q)
          if (tk[q] == key[u]) g = q;
        if (g < 0) {   // not in the snapshot: claim the first free entry, or find the key claimed meanwhile
#pragma unroll 1
          for (int q = 0; q < kTinyGroups && g < 0; ++q) {
            const unsigned long long old = atomicCAS(reinterpret_cast<unsigned long long*>(&S.keys[q]), (unsigned long long)kEmpty,
                                                     (unsigned long long)key[u]);
            if (old == kEmpty || old == key[u]) g = q;
          }
#pragma unroll
          for (int q = 0; q < kTinyGroups; ++q) tk[q] = uint32_t(*reinterpret_cast<const volatile uint64_t*>(&S.keys[q]));
        }
        if (g < 0) {   // a 9th group in this CTA: the generic kernel takes the row
          if (on) {
            atomicOr(a.defer_out + (row >> 5), 1u << (row & 31));
            atomicAdd(&S.deferred, 1u);
          }
        } else {
          if (on && !((seen >> g) & 1u)) {
            seen |= 1u << g;
            atomicMin(&S.first[g], uint32_t(row - begin));   // rows of a thread come in increasing order
          }
          if constexpr (kSelect) {
            if (on && !valid[u]) atomicAdd(&S.nulls[g], 1u);
            const int ga = (on && valid[u]) ? g : -1;
#pragma unroll
            for (int q = 0; q < kTinyGroups; ++q) {
              if (ga == q) {
                if (kFloatSum) acc[q] = __double_as_longlong(__longlong_as_double(acc[q]) + __longlong_as_double(v[u]));
                else acc[q] += v[u];
                cnt[q] += 1;
              }
            }
          } else if (valid[u]) {
#pragma unroll
            for (int q = 0; q < kTinyGroups; ++q) {
              if (g == q) {
                if (kFloatSum) acc[q] = __double_as_longlong(__longlong_as_double(acc[q]) + __longlong_as_double(v[u]));
                else acc[q] += v[u];
                cnt[q] += 1;
              }
            }
          } else {
            atomicAdd(&S.nulls[g], 1u);
          }
        }
      }
    }
  }
#pragma unroll
  for (int q = 0; q < kTinyGroups; ++q) {
    S.sum[q][t] = acc[q];
    S.cnt[q][t] = cnt[q];
  }
  __syncthreads();
  tiny_fold(a, S, kFloatSum ? kFloat : kSigned, begin);
}

#include "groupby_q1.cuh"
#include "groupby_wide.cuh"

// ---- table maintenance ----------------------------------------------------------------------
// moves every occupied slot of `src` (stride bytes each) into table `dst`; when `merge` the states
// are folded (duplicates allowed), otherwise copied (rehash: keys are unique)
__global__ void gb_move_kernel(GbSchema s, const char* src, uint64_t n_src, int src_stride, uint64_t src_cap,
                               GbTable dst, int merge, uint32_t* counters) {
  const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n_src) return;
  const char* ss = src + i * src_stride;
  const uint64_t key = *reinterpret_cast<const uint64_t*>(ss);
  int special = 0;
  if (src_cap != 0 && i >= src_cap) {  // special slots of a source table
    if (key == kEmpty) return;
    special = int(i - src_cap) + 1;
  } else if (key == kEmpty) {
    return;
  }
  // flat exports carry the special tag in first_row's place? no: they carry it in the key order —
  // exports append the two special slots last and pass src_cap = number of ordinary slots
  bool claimed;
  uint64_t slot_index;
  char* ds = global_slot(dst, s.stride, key, special, true, &claimed, &slot_index);
  if (!ds) {
    atomicExch(counters + 3, 3u);
    return;
  }
  if (claimed) atomicAdd(counters + 0, 1u);
  if (merge) {
    merge_state(ds, ss, s);
  } else {
    const uint64_t* sw = reinterpret_cast<const uint64_t*>(ss);
    uint64_t* dw = reinterpret_cast<uint64_t*>(ds);
    for (int w = 1; w < s.stride / 8; ++w) dw[w] = sw[w];
  }
}

// occupied slots -> (first_row, slot index) pairs, unordered
__global__ void gb_collect_kernel(GbTable t, int stride, uint64_t* first_rows, uint64_t* slot_idx, uint32_t* counter,
                                  const uint32_t* slot_ids, int only_unassigned) {
  const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  bool take = i < t.cap + 2;
  uint64_t first = 0;
  if (take) {
    const ulonglong2 kw = *reinterpret_cast<const ulonglong2*>(t.slots + i * stride);   // {key, first_row}: one 16-byte load
    first = kw.y;
    take = kw.x != kEmpty;
    if (take && only_unassigned && slot_ids[i] != kNoId) take = false;
  }
  // one position claim per warp (10^8 groups bumping one counter would serialise in L2)
  const unsigned m = __ballot_sync(0xffffffffu, take);
  if (m == 0) return;
  const int lane = threadIdx.x & 31;
  uint32_t base = 0;
  if (lane == __ffs(m) - 1) base = atomicAdd(counter, uint32_t(__popc(m)));
  base = __shfl_sync(0xffffffffu, base, __ffs(m) - 1);
  if (take) {
    const uint32_t pos = base + __popc(m & ((1u << lane) - 1u));
    first_rows[pos] = first;
    slot_idx[pos] = i;
  }
}

// the groups of a pending image list (see b200_groupby::pending): every image is a group
// first rows in output order through the order finalize left behind
__global__ void gb_first_rows_kernel(const char* slots, int stride, const uint64_t* order, uint32_t n, uint64_t* out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = reinterpret_cast<const uint64_t*>(slots + order[i] * uint64_t(stride))[1];
}

__global__ void gb_collect_images_kernel(const char* images, int stride, uint32_t n, uint64_t* first_rows, uint64_t* slot_idx) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    first_rows[i] = reinterpret_cast<const uint64_t*>(images + size_t(i) * stride)[1];
    slot_idx[i] = i;
  }
}

// grouper: first_row of the newly claimed slots
__global__ void gb_claimed_rows_kernel(GbTable t, int stride, const uint32_t* claimed, uint32_t n, uint64_t* first_rows,
                                       uint64_t* slot_idx) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t s = claimed[i];
  first_rows[i] = reinterpret_cast<const uint64_t*>(t.slots + s * stride)[1];
  slot_idx[i] = s;
}

__global__ void gb_assign_ids_kernel(const uint64_t* sorted_slots, uint32_t n, uint32_t base, uint32_t* slot_ids) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) slot_ids[sorted_slots[i]] = base + i;
}

__global__ void gb_fill_u32_kernel(uint32_t* p, uint64_t n, uint32_t v) {
  const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

// export: occupied slots copied densely (ordinary slots in any order; special slots tagged by key
// words that cannot occur otherwise are re-detected on import through `special_tags`)
__global__ void gb_export_kernel(GbTable t, int stride, const uint64_t* slot_idx, uint32_t n, char* out, uint8_t* tags) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t s = slot_idx[i];
  const uint64_t* sw = reinterpret_cast<const uint64_t*>(t.slots + s * stride);
  uint64_t* dw = reinterpret_cast<uint64_t*>(out + uint64_t(i) * stride);
  for (int w = 0; w < stride / 8; ++w) dw[w] = sw[w];
  // the trailing 8 bytes of slot 0..: first word stays the key; special slots are tagged in-band by
  // storing their kind in the (otherwise unused) padding word at the end of the slot
  (void)tags;
  if (s >= t.cap) dw[stride / 8 - 1] = 0xB200000000000000ull | (s - t.cap + 1);
}

__global__ void gb_import_kernel(GbSchema s, const char* src, uint32_t n, GbTable dst, uint32_t* counters) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const char* ss = src + uint64_t(i) * s.stride;
  const uint64_t* sw = reinterpret_cast<const uint64_t*>(ss);
  int special = 0;
  const uint64_t tag = sw[s.stride / 8 - 1];
  if ((tag >> 56) == 0xB2) special = int(tag & 3);
  bool claimed;
  uint64_t slot_index;
  char* ds = global_slot(dst, s.stride, sw[0], special, true, &claimed, &slot_index);
  if (!ds) {
    atomicExch(counters + 3, 4u);
    return;
  }
  if (claimed) atomicAdd(counters + 0, 1u);
  merge_state(ds, ss, s);
}

// ---- finalize -------------------------------------------------------------------------------
struct EmitOut {
  void* data;
  uint32_t* validity;  // nullable
  int32_t type;
};
struct EmitArgs {
  GbSchema s;
  GbTable t;
  const uint64_t* order;  // slot index per output rank (nullptr: dense mode, slot = rank)
  uint32_t n_groups;
  EmitOut keys[kMaxKeys];
  EmitOut aggs[2 * kMaxAggs];
  uint32_t* null_counts;  // [kMaxKeys + 2*kMaxAggs]
};

__device__ __forceinline__ void store_typed(void* base, int type, uint32_t i, uint64_t bits_signed_or_unsigned, double f) {
  switch (type) {
    case B200_UINT8: case B200_INT8: reinterpret_cast<uint8_t*>(base)[i] = uint8_t(bits_signed_or_unsigned); break;
    case B200_UINT16: case B200_INT16: reinterpret_cast<uint16_t*>(base)[i] = uint16_t(bits_signed_or_unsigned); break;
    case B200_UINT32: case B200_INT32: reinterpret_cast<uint32_t*>(base)[i] = uint32_t(bits_signed_or_unsigned); break;
    case B200_FLOAT: reinterpret_cast<float*>(base)[i] = float(f); break;
    case B200_DOUBLE: reinterpret_cast<double*>(base)[i] = f; break;
    default: reinterpret_cast<uint64_t*>(base)[i] = bits_signed_or_unsigned; break;
  }
}

__device__ __forceinline__ uint64_t anti_extreme_bits(int type, bool for_min) {
  // Extrema / numeric_limits (hash_aggregate.cc:627-640): min starts at max(), max at min()
  switch (type) {
    case B200_INT8: return for_min ? 0x7Full : 0x80ull;
    case B200_INT16: return for_min ? 0x7FFFull : 0x8000ull;
    case B200_INT32: return for_min ? 0x7FFFFFFFull : 0x80000000ull;
    case B200_INT64: return for_min ? 0x7FFFFFFFFFFFFFFFull : 0x8000000000000000ull;
    default: return for_min ? ~0ull : 0ull;  // unsigned
  }
}

__global__ void gb_emit_kernel(const __grid_constant__ EmitArgs e) {
  const uint32_t r = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31;
  const bool in = r < e.n_groups;
  const uint64_t slot_index = in ? (e.order ? e.order[r] : r) : 0;
  const char* slot = e.t.slots + slot_index * e.s.stride;
  const uint64_t* w = reinterpret_cast<const uint64_t*>(slot);
  // The slots are visited in first-row order, i.e. at random: every load of a slot is a DRAM round trip, and the
  // decode below (run-time aggregate kinds, warp votes in between) would take them one after the other.  The first
  // 64 bytes of the slot are requested up front, all loads in flight together; states beyond them are read in place.
  constexpr int kPre = 8;
  uint64_t pre[kPre];
  const int slot_words = e.s.stride / 8;
#pragma unroll
  for (int i = 0; i < kPre; ++i) pre[i] = i < slot_words ? w[i] : 0;
  auto word = [&](const uint64_t* st, int i) -> uint64_t {
    const int idx = int(st - w) + i;
    uint64_t v = pre[0];
#pragma unroll
    for (int q = 1; q < kPre; ++q) v = idx == q ? pre[q] : v;
    return idx < kPre ? v : st[i];
  };
  const uint64_t key = pre[0];
  const bool special_null = e.order && e.s.null_shift < 0 && slot_index == e.t.cap;
  const bool special_full = e.order && e.s.null_shift < 0 && slot_index == e.t.cap + 1;
  // keys: GetUniques / Decode (hash_aggregate.cc:412-429, 65-94, 174-192)
  for (int k = 0; k < e.s.n_keys; ++k) {
    const EmitOut& o = e.keys[k];
    if (!o.data) continue;
    bool valid = true;
    uint64_t v = 0;
    if (in) {
      if (e.s.null_shift >= 0) valid = !((key >> (e.s.null_shift + k)) & 1ull);
      else valid = !special_null;
      const int nb = e.s.key_bits[k];
      if (valid) {
        const uint64_t kk = special_full ? kEmpty : key;
        v = nb == 64 ? kk : ((kk >> e.s.key_shift[k]) & ((uint64_t(1) << nb) - 1));
      }
    }
    if (e.s.key_type[k] == B200_BOOL) {
      const unsigned word = __ballot_sync(0xffffffffu, in && valid && v);
      if (lane == 0 && (r >> 5) <= ((e.n_groups - 1) >> 5)) reinterpret_cast<uint32_t*>(o.data)[r >> 5] = word;
    } else if (in) {
      // raw key bits (floats group by bit pattern)
      switch (e.s.key_bits[k]) {
        case 8: reinterpret_cast<uint8_t*>(o.data)[r] = uint8_t(v); break;
        case 16: reinterpret_cast<uint16_t*>(o.data)[r] = uint16_t(v); break;
        case 32: reinterpret_cast<uint32_t*>(o.data)[r] = uint32_t(v); break;
        default: reinterpret_cast<uint64_t*>(o.data)[r] = v; break;
      }
    }
    const unsigned vw = __ballot_sync(0xffffffffu, in && valid);
    const unsigned inw = __ballot_sync(0xffffffffu, in);
    if (lane == 0 && inw) {
      if (o.validity) o.validity[r >> 5] = vw;
      const int nulls = __popc(inw & ~vw);
      if (nulls) atomicAdd(e.null_counts + k, uint32_t(nulls));
    }
  }
  int oi = 0;
  for (int j = 0; j < e.s.n_aggs; ++j) {
    const GbAgg& g = e.s.aggs[j];
    const uint64_t* st = reinterpret_cast<const uint64_t*>(slot + g.off);
    bool valid = true;
    if (g.kind == B200_HASH_MIN_MAX) {
      const EmitOut& omin = e.aggs[oi];
      const EmitOut& omax = e.aggs[oi + 1];
      if (in) {
        const uint64_t flags = word(st, 3);
        const bool has_values = flags & 1, has_nulls = flags & 2;
        valid = has_values && !(g.option == 1 && has_nulls);  // hash_aggregate.cc:722-732
        if (g.dom == kFloat) {
          double mn = __longlong_as_double(0x7FF0000000000000ull), mx = __longlong_as_double(0xFFF0000000000000ull);
          if (word(st, 0) != ~uint64_t(0)) {  // at least one non-NaN value
            mn = unord_f64(word(st, 0));
            mx = unord_f64(word(st, 1));
            if (word(st, 2) != ~uint64_t(0)) {  // a zero was seen: the earliest zero row decides the sign
              const double z = (word(st, 2) & 1ull) ? -0.0 : 0.0;
              if (mn == 0.0) mn = z;
              if (mx == 0.0) mx = z;
            }
          }
          store_typed(omin.data, g.arg_type, r, 0, mn);
          store_typed(omax.data, g.arg_type, r, 0, mx);
        } else {
          uint64_t mn = anti_extreme_bits(g.arg_type, true), mx = anti_extreme_bits(g.arg_type, false);
          if (has_values) {
            mn = g.dom == kSigned ? (word(st, 0) ^ 0x8000000000000000ull) : word(st, 0);
            mx = g.dom == kSigned ? (word(st, 1) ^ 0x8000000000000000ull) : word(st, 1);
          }
          store_typed(omin.data, g.arg_type, r, mn, 0.0);
          store_typed(omax.data, g.arg_type, r, mx, 0.0);
        }
      }
      const unsigned vw = __ballot_sync(0xffffffffu, in && valid);
      const unsigned inw = __ballot_sync(0xffffffffu, in);
      if (lane == 0 && inw) {
        if (omin.validity) omin.validity[r >> 5] = vw;
        if (omax.validity && omax.validity != omin.validity) omax.validity[r >> 5] = vw;
        const int nulls = __popc(inw & ~vw);
        if (nulls) {
          atomicAdd(e.null_counts + kMaxKeys + oi, uint32_t(nulls));
          atomicAdd(e.null_counts + kMaxKeys + oi + 1, uint32_t(nulls));
        }
      }
      oi += 2;
      continue;
    }
    const EmitOut& o = e.aggs[oi];
    if (in) {
      if (g.kind == B200_HASH_COUNT) {
        reinterpret_cast<int64_t*>(o.data)[r] = int64_t(word(st, 0));
      } else if (g.kind == B200_HASH_SUM) {
        valid = word(st, 1) > 0;  // hash_aggregate.cc:595-605
        reinterpret_cast<uint64_t*>(o.data)[r] = word(st, 0);
      } else {              // MeanImpl::Finalize: static_cast<double>(sum) / count
        valid = word(st, 1) > 0;
        double s = g.dom == kFloat ? __longlong_as_double(word(st, 0))
                                   : (g.dom == kSigned ? double(int64_t(word(st, 0))) : double(word(st, 0)));
        reinterpret_cast<double*>(o.data)[r] = valid ? s / double(word(st, 1)) : 0.0;
      }
    }
    const unsigned vw = __ballot_sync(0xffffffffu, in && valid);
    const unsigned inw = __ballot_sync(0xffffffffu, in);
    if (lane == 0 && inw) {
      if (o.validity) o.validity[r >> 5] = vw;
      const int nulls = __popc(inw & ~vw);
      if (nulls) atomicAdd(e.null_counts + kMaxKeys + oi, uint32_t(nulls));
    }
    oi += 1;
  }
}

// ---- dense (group-id indexed) accumulation: HashAggregateKernel::consume ----------------------
__global__ void gb_dense_consume_kernel(GbSchema s, int agg_index, GbCol col, const uint32_t* ids, int64_t n,
                                        char* dense, uint32_t num_groups, uint32_t* counters) {
  const int64_t row = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (row >= n) return;
  const uint32_t g = ids[row];
  if (g >= num_groups) {
    atomicExch(counters + 3, 5u);
    return;
  }
  const Val v = load_val(col, row);
  accumulate(dense + uint64_t(g) * s.stride, s.aggs[agg_index], v, uint64_t(row));
}

// ---- hash repartitioning (multi-GPU, high cardinality: SURVEY §8e) ---------------------------------
// dest(row) = (mix64(packed key) >> 40) % n_parts — bits disjoint from the table's slot index, so the
// keys one rank receives still spread over its table.  Rows of one destination end up contiguous.
constexpr int kPartMax = 16;
constexpr int kPartThreads = 256;
constexpr int kPartItems = 16;
constexpr int kPartTile = kPartThreads * kPartItems;

__device__ __forceinline__ uint32_t part_of(const GbArgs& a, int64_t row, uint32_t n_parts) {
  int special;
  const uint64_t key = load_key(a, row, &special);
  if (special) return 0;   // the null key / the EMPTY-valued key: one fixed owner
  return uint32_t((mix64(key) >> 40) % n_parts);
}

// pass 1: per-tile destination counts
__global__ void __launch_bounds__(kPartThreads) gb_part_count_kernel(const __grid_constant__ GbArgs a, uint32_t n_parts,
                                                                     uint32_t* tile_counts) {
  __shared__ uint32_t cnt[kPartMax];
  if (threadIdx.x < kPartMax) cnt[threadIdx.x] = 0;
  __syncthreads();
  const int64_t base = int64_t(blockIdx.x) * kPartTile;
#pragma unroll 4
  for (int i = 0; i < kPartItems; ++i) {
    const int64_t row = base + i * kPartThreads + threadIdx.x;
    if (row < a.n) atomicAdd(&cnt[part_of(a, row, n_parts)], 1u);
  }
  __syncthreads();
  if (threadIdx.x < n_parts) tile_counts[size_t(blockIdx.x) * kPartMax + threadIdx.x] = cnt[threadIdx.x];
}

// pass 2 (one CTA): per destination, exclusive scan over tiles; totals and destination bases
__global__ void __launch_bounds__(1024) gb_part_scan_kernel(uint32_t* tile_counts, int64_t n_tiles, uint32_t n_parts,
                                                            int64_t* totals /* [n_parts] counts, [kPartMax + p] bases */) {
  __shared__ uint32_t warp_sums[32];
  __shared__ uint32_t carry_s;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  for (uint32_t p = 0; p < n_parts; ++p) {
    if (t == 0) carry_s = 0;
    __syncthreads();
    for (int64_t base = 0; base < n_tiles; base += 1024) {
      const int64_t tile = base + t;
      const uint32_t c = tile < n_tiles ? tile_counts[size_t(tile) * kPartMax + p] : 0;
      uint32_t incl = c;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
      }
      if (lane == 31) warp_sums[warp] = incl;
      __syncthreads();
      uint32_t wpre = 0;
      for (int w = 0; w < warp; ++w) wpre += warp_sums[w];
      const uint32_t carry = carry_s;
      if (tile < n_tiles) tile_counts[size_t(tile) * kPartMax + p] = carry + wpre + incl - c;
      __syncthreads();
      if (t == 1023) carry_s = carry + wpre + incl;
      __syncthreads();
    }
    if (t == 0) totals[p] = carry_s;
    __syncthreads();
  }
  if (t == 0) {
    int64_t acc = 0;
    for (uint32_t p = 0; p < n_parts; ++p) {
      totals[kPartMax + p] = acc;
      acc += totals[p];
    }
  }
}

// pass 3: position of every row in the partitioned layout
__global__ void __launch_bounds__(kPartThreads) gb_part_pos_kernel(const __grid_constant__ GbArgs a, uint32_t n_parts,
                                                                   const uint32_t* tile_offsets, const int64_t* totals,
                                                                   uint32_t* pos) {
  __shared__ uint32_t cur[kPartMax];
  if (threadIdx.x < kPartMax)
    cur[threadIdx.x] = threadIdx.x < n_parts ? uint32_t(totals[kPartMax + threadIdx.x]) + tile_offsets[size_t(blockIdx.x) * kPartMax + threadIdx.x] : 0;
  __syncthreads();
  const int64_t base = int64_t(blockIdx.x) * kPartTile;
#pragma unroll 4
  for (int i = 0; i < kPartItems; ++i) {
    const int64_t row = base + i * kPartThreads + threadIdx.x;
    if (row < a.n) pos[row] = atomicAdd(&cur[part_of(a, row, n_parts)], 1u);
  }
}

template <typename T>
__global__ void gb_part_scatter_kernel(const T* __restrict__ in, BitmapRef vv, const uint32_t* __restrict__ pos, int64_t n,
                                       T* __restrict__ out, uint8_t* __restrict__ out_valid) {
  const int64_t row = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (row >= n) return;
  const uint32_t p = pos[row];
  out[p] = in[row];
  if (out_valid) out_valid[p] = bitmap_get(vv, row) ? 1 : 0;
}

__global__ void gb_part_rows_kernel(int64_t n, uint64_t row_base, const uint32_t* __restrict__ pos, uint64_t* __restrict__ out) {
  const int64_t row = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (row < n) out[pos[row]] = row_base + uint64_t(row);
}

// validity bytes (0/1 per row) -> LSB-first bitmap, whole 32-bit words; counts nulls
__global__ void gb_pack_validity_kernel(const uint8_t* __restrict__ bytes, int64_t n, uint32_t* __restrict__ bitmap,
                                        unsigned long long* null_count) {
  const int64_t row = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  const bool in = row < n;
  const bool v = in && bytes[row] != 0;
  const unsigned word = __ballot_sync(0xffffffffu, v);
  const unsigned inw = __ballot_sync(0xffffffffu, in);
  if ((threadIdx.x & 31) == 0 && inw) {
    bitmap[row >> 5] = word;
    const int nulls = __popc(inw & ~word);
    if (nulls) atomicAdd(null_count, (unsigned long long)nulls);
  }
}

// dense states over one id space folded slot by slot: HashAggregateKernel::merge
__global__ void gb_dense_merge_kernel(GbSchema s, const char* src, char* dst, uint32_t n) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) merge_state(dst + uint64_t(i) * s.stride, src + uint64_t(i) * s.stride, s);
}

}  // namespace

// ===============================================================================================
struct b200_groupby {
  b200_ctx* ctx = nullptr;
  GbSchema schema;
  std::vector<b200_agg_spec> specs;
  GbTable table{nullptr, 0};
  uint32_t* counters = nullptr;  // device, 8 words
  uint32_t groups = 0;           // host view, refreshed at every sync point
  uint64_t expected = 0;         // caller's cardinality hint
  bool fast_disabled = false;    // observed cardinality exceeds the shared-memory table
  uint32_t spilled = 0;          // fast path: groups waiting in the spill list
  // grouper state
  uint32_t* slot_ids = nullptr;
  uint64_t slot_ids_cap = 0;
  uint32_t ids_assigned = 0;
  // dense mode (b200_hash_agg_consume)
  char* dense = nullptr;
  uint32_t dense_groups = 0;
  uint32_t dense_cap = 0;
  bool dense_used = false;
  uint64_t dense_rows = 0;
  // scratch owned by the handle
  void* tmp = nullptr;
  size_t tmp_bytes = 0;
  // Groups of a first, fully partitioned consume wait here as slot images (inside hp_tmp) instead of in the table:
  // finalize reads them directly; any other use of the handle folds them into the table first (materialize()).
  // With a cardinality hint in the partitioned path's range the table itself is only allocated when needed:
  // at 10^8 expected groups it is a 17 GB array to initialise, fill and scan.
  char* pending = nullptr;
  uint32_t pending_n = 0;
  // wide keys (groupby_wide.cuh): the key columns as the caller declared them, the key -> id table and the key store;
  // `schema` then describes ONE uint32 key column holding the ids
  bool wide = false;
  int32_t wn_keys = 0, wwords = 0;
  int32_t wtypes[kWideMaxKeys], wbits[kWideMaxKeys], woff[kWideMaxKeys];
  uint64_t* wslots = nullptr;
  uint64_t wcap = 0;
  uint64_t* wstore = nullptr;
  uint32_t wstore_cap = 0, wcount = 0;
  uint32_t* wcounters = nullptr;
  void* wtmp = nullptr;
  size_t wtmp_bytes = 0;
  // variable-length key columns: heap of the stored keys' bytes, where each stored key's bytes start, bytes per column
  int32_t wn_var = 0;
  int32_t wvar_col[kWideMaxKeys], wvar_index[kWideMaxKeys];
  uint8_t* wheap = nullptr;
  uint64_t wheap_cap = 0, wheap_used = 0;
  uint64_t* wstore_var = nullptr;
  unsigned long long* wvar_bytes = nullptr;   // device [kWideMaxKeys]
  uint64_t wvar_bytes_host[kWideMaxKeys] = {};
  // the group order finalize computed (slot / image index per output rank, inside `tmp`): b200_groupby_first_rows right
  // after a finalize gathers through it instead of collecting and sorting once more; dropped by any change of state
  int64_t hp_calls = 0, hp_alloc_failures = 0, hp_span_refusals = 0;   // b200_groupby_stat
  const uint64_t* order_cache = nullptr;
  const char* order_slots = nullptr;
  uint32_t order_n = 0;
  // upper bound of every global row number consumed so far (2^64-1 once unknown): first rows below 2^32 are ranked
  // by a sort that carries 32-bit keys
  uint64_t row_bound = 0;
  // high-cardinality path (groupby_hp.cuh): partition buffers, and the sizes from which it is taken
  void* hp_tmp = nullptr;
  size_t hp_bytes = 0;
  int64_t hp_min_rows = int64_t(1) << 22;
  int64_t hp_min_groups = int64_t(1) << 21;   // a table of this many slots no longer fits in L2
  int64_t hp_keys_per_part = kHpKeysPerPart;  // target distinct keys per final partition
};

namespace {

int domain_of(int t) {
  switch (t) {
    case B200_FLOAT: case B200_DOUBLE: return kFloat;
    case B200_INT8: case B200_INT16: case B200_INT32: case B200_INT64: return kSigned;
    default: return kUnsigned;
  }
}

int gb_reserve_tmp(b200_groupby* gb, size_t bytes) {
  b200_ctx* ctx = gb->ctx;
  gb->order_cache = nullptr;   // whoever asks for scratch is about to overwrite it
  if (bytes <= gb->tmp_bytes) return B200_OK;
  if (gb->tmp) {
    b200_dev_free(ctx, gb->tmp);
    gb->tmp = nullptr;
    gb->tmp_bytes = 0;
  }
  bytes = b200_align_up(bytes, 1 << 20);
  B200_RETURN_NOT_OK(b200_dev_alloc(ctx, &gb->tmp, bytes));
  gb->tmp_bytes = bytes;
  return B200_OK;
}

int alloc_table(b200_groupby* gb, uint64_t cap, GbTable* out) {
  b200_ctx* ctx = gb->ctx;
  GbTable t;
  t.cap = cap;
  const uint64_t n_slots = cap + 2;
  B200_RETURN_NOT_OK(b200_dev_alloc(ctx, reinterpret_cast<void**>(&t.slots), n_slots * gb->schema.stride));
  gb_init_kernel<<<(unsigned)((n_slots + 255) / 256), 256, 0, ctx->stream>>>(gb->schema, t.slots, n_slots, gb->schema.stride);
  B200_LAUNCHED(ctx);
  *out = t;
  return B200_OK;
}

uint64_t pow2_at_least(uint64_t x) {
  uint64_t p = 1024;
  while (p < x) p <<= 1;
  return p;
}

int read_counters(b200_groupby* gb, uint32_t out[4]) {
  b200_ctx* ctx = gb->ctx;
  B200_CUDA(ctx, cudaMemcpyAsync(ctx->host_words + 16, gb->counters, 8 * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
  B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  memcpy(out, ctx->host_words + 16, 4 * sizeof(uint32_t));
  gb->groups = out[0];
  gb->spilled = reinterpret_cast<const uint32_t*>(ctx->host_words + 16)[5];
  if (out[3] != 0) return b200_fail(ctx, B200_UNKNOWN_ERROR, "groupby: internal table error %u", out[3]);
  return B200_OK;
}

// grows the table to at least `min_cap` slots, re-inserting every group
uint64_t initial_cap(const b200_groupby* gb) { return pow2_at_least(gb->expected > 0 ? gb->expected * 2 : (1u << 16)); }

int grow_table(b200_groupby* gb, uint64_t min_cap) {
  b200_ctx* ctx = gb->ctx;
  gb->order_cache = nullptr;
  if (!gb->table.slots) return alloc_table(gb, std::max(pow2_at_least(min_cap), initial_cap(gb)), &gb->table);
  const uint64_t new_cap = pow2_at_least(min_cap);
  if (new_cap <= gb->table.cap) return B200_OK;
  GbTable nt;
  B200_RETURN_NOT_OK(alloc_table(gb, new_cap, &nt));
  // group counter is rebuilt by the move
  B200_CUDA(ctx, cudaMemsetAsync(gb->counters, 0, sizeof(uint32_t), ctx->stream));
  const uint64_t n_src = gb->table.cap + 2;
  gb_move_kernel<<<(unsigned)((n_src + 255) / 256), 256, 0, ctx->stream>>>(gb->schema, gb->table.slots, n_src,
                                                                           gb->schema.stride, gb->table.cap, nt, 0,
                                                                           gb->counters);
  B200_LAUNCHED(ctx);
  B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  b200_dev_free(ctx, gb->table.slots);
  gb->table = nt;
  // slot ids are positional: the grouper re-derives them after a rehash
  if (gb->slot_ids) {
    b200_dev_free(ctx, gb->slot_ids);
    gb->slot_ids = nullptr;
    gb->slot_ids_cap = 0;
    gb->ids_assigned = 0;
  }
  return B200_OK;
}

int ensure_table(b200_groupby* gb) { return gb->table.slots ? B200_OK : alloc_table(gb, initial_cap(gb), &gb->table); }

// the table exists and holds every group (pending images folded in)
int materialize(b200_groupby* gb) {
  b200_ctx* ctx = gb->ctx;
  if (gb->pending) gb->order_cache = nullptr;
  B200_RETURN_NOT_OK(ensure_table(gb));
  if (!gb->pending) return B200_OK;
  const uint32_t n = gb->pending_n;
  const char* images = gb->pending;
  gb->pending = nullptr;
  gb->pending_n = 0;
  gb->groups = 0;
  B200_CUDA(ctx, cudaMemsetAsync(gb->counters, 0, sizeof(uint32_t), ctx->stream));
  B200_RETURN_NOT_OK(grow_table(gb, 2 * uint64_t(n)));
  B200_TIMED(ctx, "groupby_import", (gb_hp_import_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(gb->schema, images, n, gb->table, gb->counters)));
  uint32_t c[4];
  return read_counters(gb, c);
}

int fill_keys(b200_groupby* gb, const b200_col* keys, GbArgs* a) {
  b200_ctx* ctx = gb->ctx;
  const int64_t n = keys[0].length;
  for (int k = 0; k < gb->schema.n_keys; ++k) {
    const b200_col& c = keys[k];
    if (c.type != gb->schema.key_type[k]) return b200_fail(ctx, B200_TYPE_ERROR, "groupby: key %d has type %d, expected %d", k, c.type, gb->schema.key_type[k]);
    if (c.length != n) return b200_fail(ctx, B200_INVALID, "Array arguments must all be the same length");
    if (n > 0 && !c.data) return b200_fail(ctx, B200_INVALID, "groupby: key %d has no data buffer", k);
    GbKey& g = a->keys[k];
    const int bits = gb->schema.key_bits[k];
    g.type = c.type;
    g.nbits = bits;
    g.shift = gb->schema.key_shift[k];
    g.data = bits >= 8 ? reinterpret_cast<const char*>(c.data) + c.offset * (bits / 8) : nullptr;
    g.bits = make_bitmap_ref(bits == 1 ? c.data : nullptr, c.offset);
    const bool has_nulls = c.validity != nullptr && c.null_count != 0;
    g.vv = has_nulls ? make_bitmap_ref(c.validity, c.offset) : BitmapRef{nullptr, 0};
  }
  a->n = n;
  return B200_OK;
}

void fill_col(const b200_col& c, GbCol* g) {
  const int bits = b200_type_bits(c.type);
  g->type = c.type;
  g->data = bits >= 8 ? reinterpret_cast<const char*>(c.data) + c.offset * (bits / 8) : nullptr;
  g->bits = make_bitmap_ref(bits == 1 ? c.data : nullptr, c.offset);
  const bool has_nulls = c.validity != nullptr && c.null_count != 0;
  g->vv = has_nulls ? make_bitmap_ref(c.validity, c.offset) : BitmapRef{nullptr, 0};
}

// groups a CTA could not place in the global table wait as slot images in the spill list
int import_spilled(b200_groupby* gb, const char* spill) {
  b200_ctx* ctx = gb->ctx;
  if (!gb->spilled) return B200_OK;
  const uint32_t n_spill = gb->spilled;
  B200_RETURN_NOT_OK(grow_table(gb, 2 * (uint64_t(gb->groups) + n_spill)));
  gb_import_kernel<<<(n_spill + 255) / 256, 256, 0, ctx->stream>>>(gb->schema, spill, n_spill, gb->table, gb->counters);
  B200_LAUNCHED(ctx);
  B200_CUDA(ctx, cudaMemsetAsync(gb->counters + 5, 0, sizeof(uint32_t), ctx->stream));
  uint32_t c[4];
  return read_counters(gb, c);
}


// ---- high-cardinality path: partition by hash, aggregate the partitions in shared memory, import the groups
bool hp_applicable(const b200_groupby* gb, const GbArgs& a, uint64_t likely_groups) {
  if (a.mode != kModeFused || a.n_cols > 1 || a.defer_in) return false;
  if (a.n < gb->hp_min_rows || a.n >= (int64_t(1) << 31)) return false;
  if (likely_groups < uint64_t(gb->hp_min_groups)) return false;
  for (int j = 0; j < a.s.n_aggs; ++j) {
    const int kind = a.s.aggs[j].kind;
    if (kind != B200_HASH_COUNT && kind != B200_HASH_SUM && kind != B200_HASH_MEAN) return false;
  }
  return true;
}

// *handled = false: not taken (explicit row numbers out of order); the caller goes on with the generic kernels
int run_hp_consume(b200_groupby* gb, GbArgs& a, uint64_t likely_groups, bool* handled, uint32_t* deferred) {
  b200_ctx* ctx = gb->ctx;
  *handled = false;
  *deferred = 0;
  const int64_t n = a.n;
  // partition plan
  const uint64_t est = std::min<uint64_t>(uint64_t(n), std::max<uint64_t>(likely_groups, 1));
  int bits = 1;
  while (bits < 17 && (uint64_t(gb->hp_keys_per_part) << bits) < est) ++bits;
  const int bits1 = bits <= 9 ? bits : bits / 2;
  const int bits2 = bits - bits1;
  const uint32_t parts1 = 1u << bits1, parts = 1u << bits;
  // scratch: two tuple buffers + histograms / cursors / starts + work items
  const size_t a8 = b200_align_up(size_t(n) * 8, 256), a4 = b200_align_up(size_t(n) * 4, 256);
  const size_t buf_bytes = 2 * a8 + a4 + (a.row_ids ? a4 : 0);
  const size_t small = b200_align_up(size_t(parts + 1) * 4, 256);
  const size_t need = 2 * buf_bytes + 9 * small + 1024;
  if (need > gb->hp_bytes) {
    if (gb->hp_tmp) b200_dev_free(ctx, gb->hp_tmp);
    gb->hp_tmp = nullptr;
    gb->hp_bytes = 0;
    if (b200_dev_alloc(ctx, &gb->hp_tmp, need) != B200_OK) {
      ctx->error.clear();
      ++gb->hp_alloc_failures;
      return B200_OK;   // not enough memory for the partition buffers: the generic path needs none
    }
    gb->hp_bytes = need;
  }
  char* base = reinterpret_cast<char*>(gb->hp_tmp);
  auto bufs_at = [&](char* p) {
    HpBufs b;
    b.keys = reinterpret_cast<uint64_t*>(p);
    b.vals = reinterpret_cast<uint64_t*>(p + a8);
    b.rows = reinterpret_cast<uint32_t*>(p + 2 * a8);
    b.ord = a.row_ids ? reinterpret_cast<uint32_t*>(p + 2 * a8 + a4) : nullptr;
    return b;
  };
  const HpBufs buf_a = bufs_at(base), buf_b = bufs_at(base + buf_bytes);
  char* sm = base + 2 * buf_bytes;
  uint32_t* hist1 = reinterpret_cast<uint32_t*>(sm);
  uint32_t* hist2 = reinterpret_cast<uint32_t*>(sm + small);
  uint32_t* items = reinterpret_cast<uint32_t*>(sm + 2 * small);
  uint32_t* misc = reinterpret_cast<uint32_t*>(sm + 3 * small);   // [0] ticket, [1] row-order flag
  const size_t zero_bytes = 3 * small + 256;
  uint32_t* cursor1 = reinterpret_cast<uint32_t*>(sm + 3 * small + 256);
  uint32_t* base1 = reinterpret_cast<uint32_t*>(sm + 4 * small + 256);
  uint32_t* cursor2 = reinterpret_cast<uint32_t*>(sm + 5 * small + 256);
  uint32_t* base2 = reinterpret_cast<uint32_t*>(sm + 6 * small + 256);
  uint32_t* item_base = reinterpret_cast<uint32_t*>(sm + 7 * small + 256);
  B200_CUDA(ctx, cudaMemsetAsync(sm, 0, zero_bytes, ctx->stream));
  uint64_t ord_base = 0;
  if (a.row_ids) {
    // explicit row numbers ride along as 32-bit offsets from the batch's smallest one (the earliest row of a group is
    // the minimum of them); a batch whose numbers span 2^31 or more goes through the generic kernels
    unsigned long long* mm = reinterpret_cast<unsigned long long*>(misc + 2);
    const unsigned long long init[2] = {~0ull, 0ull};
    B200_CUDA(ctx, cudaMemcpyAsync(mm, init, sizeof(init), cudaMemcpyHostToDevice, ctx->stream));
    gb_hp_minmax_kernel<<<ctx->num_sms * 4, 256, 0, ctx->stream>>>(a.row_ids, n, mm);
    B200_LAUNCHED(ctx);
    B200_CUDA(ctx, cudaMemcpyAsync(ctx->host_words + 24, mm, 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
    B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const uint64_t lo = uint64_t(ctx->host_words[24]), hi = uint64_t(ctx->host_words[25]);
    if (hi - lo >= (uint64_t(1) << 31)) {
      ++gb->hp_span_refusals;
      return B200_OK;
    }
    ord_base = lo;
    a.row_base = lo;   // (unused by the kernels that read row_ids) the aggregation kernel adds it to the smallest offset
  }
  // the table takes the expected groups at load <= 1/2; a partition = the hash bits just below its index width
  // (the table itself may not exist yet, see b200_groupby::pending)
  const uint64_t planned_cap = std::max<uint64_t>(gb->table.cap, pow2_at_least(2 * std::max<uint64_t>(likely_groups, 32 * 1024)));
  int cap_bits = 0;
  while ((uint64_t(1) << cap_bits) < planned_cap) ++cap_bits;
  const int top = std::max(cap_bits, bits);   // partition id = hash bits [top - bits, top)
  if (b200_ctx_once(ctx, 1)) {
    B200_CUDA(ctx, cudaFuncSetAttribute(gb_hp_part_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(offsetof(HpPartSmem, ord))));
    B200_CUDA(ctx, cudaFuncSetAttribute(gb_hp_part_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(offsetof(HpPartSmem, ord))));
    B200_CUDA(ctx, cudaFuncSetAttribute(gb_hp_part_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(HpPartSmem))));
    B200_CUDA(ctx, cudaFuncSetAttribute(gb_hp_part_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(HpPartSmem))));
    B200_CUDA(ctx, cudaFuncSetAttribute(gb_hp_agg_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(HpAggSmem))));
    B200_CUDA(ctx, cudaFuncSetAttribute(gb_hp_agg_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(HpAggSmem))));
  }
  const int64_t n_tiles = (n + kHpTile - 1) / kHpTile;
  const unsigned grid = unsigned(std::min<int64_t>(n_tiles, int64_t(ctx->num_sms) * 4 * B200_HP_PART_MINBLOCKS));
  HpArgs h;
  memset(&h, 0, sizeof(h));
  h.has_val = a.n_cols > 0;
  h.ord_base = ord_base;
  if (a.s.n_keys == 1 && !a.keys[0].vv.words && (a.keys[0].nbits == 64 || a.keys[0].nbits == 32) &&
      (a.n_cols == 0 || (b200_type_bits(a.cols[0].type) == 64 && !a.cols[0].vv.words)))
    h.fast = a.keys[0].nbits / 8;
  // level 1
  h.dst = buf_a;
  h.parts = 1;
  h.hist = hist1;
  h.cursor = cursor1;
  h.bins = int(parts1);
  h.shift = top - bits1;
  B200_TIMED(ctx, "groupby_partition", (gb_hp_hist_kernel<1><<<grid, kHpThreads, 0, ctx->stream>>>(a, h)));
  gb_hp_scan_kernel<<<1, 1024, 0, ctx->stream>>>(hist1, parts1, cursor1, base1);
  B200_LAUNCHED(ctx);
  if (a.row_ids)
    B200_TIMED(ctx, "groupby_partition", (gb_hp_part_kernel<1, true><<<grid, kHpThreads, int(sizeof(HpPartSmem)), ctx->stream>>>(a, h)));
  else
    B200_TIMED(ctx, "groupby_partition", (gb_hp_part_kernel<1, false><<<grid, kHpThreads, int(offsetof(HpPartSmem, ord)), ctx->stream>>>(a, h)));
  HpBufs agg_src = buf_a;
  const uint32_t* agg_base = base1;
  char* images = reinterpret_cast<char*>(buf_b.keys);
  if (bits2 > 0) {
    h.src = buf_a;
    h.dst = buf_b;
    h.base = base1;
    h.parts = int(parts1);
    h.hist = hist2;
    h.cursor = cursor2;
    h.bins = 1 << bits2;
    h.shift = top - bits;
    B200_TIMED(ctx, "groupby_partition", (gb_hp_hist_kernel<2><<<grid, kHpThreads, 0, ctx->stream>>>(a, h)));
    gb_hp_scan_kernel<<<1, 1024, 0, ctx->stream>>>(hist2, parts, cursor2, base2);
    B200_LAUNCHED(ctx);
    if (a.row_ids)
      B200_TIMED(ctx, "groupby_partition", (gb_hp_part_kernel<2, true><<<grid, kHpThreads, int(sizeof(HpPartSmem)), ctx->stream>>>(a, h)));
    else
      B200_TIMED(ctx, "groupby_partition", (gb_hp_part_kernel<2, false><<<grid, kHpThreads, int(offsetof(HpPartSmem, ord)), ctx->stream>>>(a, h)));
    agg_src = buf_b;
    agg_base = base2;
    images = reinterpret_cast<char*>(buf_a.keys);
  }
  // aggregation: one work item per partition (several for a partition far above the average size)
  const uint32_t seg_rows = uint32_t(std::max<int64_t>(int64_t(1) << 16, 8 * (n / parts)));
  gb_hp_items_kernel<<<(parts + 255) / 256, 256, 0, ctx->stream>>>(agg_base, parts, seg_rows, items);
  B200_LAUNCHED(ctx);
  gb_hp_scan_kernel<<<1, 1024, 0, ctx->stream>>>(items, parts, nullptr, item_base);
  B200_LAUNCHED(ctx);
  const uint32_t image_cap = uint32_t(std::min<uint64_t>(buf_bytes / size_t(a.s.stride), 0xFFFFFFF0ull));
  B200_CUDA(ctx, cudaMemsetAsync(gb->counters + 6, 0, sizeof(uint32_t), ctx->stream));
  bool float_sum = false;
  for (int j = 0; j < a.s.n_aggs; ++j)
    if (a.s.aggs[j].kind != B200_HASH_COUNT && a.s.aggs[j].dom == kFloat) float_sum = true;
  const unsigned agg_grid = unsigned(std::min<int64_t>(int64_t(parts), int64_t(ctx->num_sms) * 2));
  if (float_sum)
    B200_TIMED(ctx, "groupby_consume", (gb_hp_agg_kernel<true><<<agg_grid, kHpThreads, int(sizeof(HpAggSmem)), ctx->stream>>>(
                                           a, agg_src, agg_base, item_base, parts, seg_rows, images, image_cap, misc)));
  else
    B200_TIMED(ctx, "groupby_consume", (gb_hp_agg_kernel<false><<<agg_grid, kHpThreads, int(sizeof(HpAggSmem)), ctx->stream>>>(
                                           a, agg_src, agg_base, item_base, parts, seg_rows, images, image_cap, misc)));
  uint32_t c[4];
  B200_RETURN_NOT_OK(read_counters(gb, c));
  *deferred = c[1];
  const uint32_t n_images = std::min(reinterpret_cast<const uint32_t*>(ctx->host_words + 16)[6], image_cap);
  const uint32_t n_items = reinterpret_cast<const uint32_t*>(ctx->host_words + 16)[7];
  *handled = true;
  ++gb->hp_calls;
  if (gb->groups == 0 && c[1] == 0 && n_items <= parts && n_images > 0) {
    // first groups of the handle, every row placed, no partition cut into several work items: the images are the
    // groups, each key once.  They stay where they are until something needs the table.
    gb->pending = images;
    gb->pending_n = n_images;
    gb->groups = n_images;
    gb_fill_u32_kernel<<<1, 32, 0, ctx->stream>>>(gb->counters, 1, n_images);
    B200_LAUNCHED(ctx);
    return B200_OK;
  }
  B200_RETURN_NOT_OK(ensure_table(gb));
  if (n_images) {
    if (2 * (uint64_t(gb->groups) + n_images) > gb->table.cap) B200_RETURN_NOT_OK(grow_table(gb, 2 * (uint64_t(gb->groups) + n_images)));
    B200_TIMED(ctx, "groupby_import", (gb_hp_import_kernel<<<(n_images + 255) / 256, 256, 0, ctx->stream>>>(gb->schema, images, n_images,
                                                                                                           gb->table, gb->counters)));
    B200_RETURN_NOT_OK(read_counters(gb, c));
  }
  return B200_OK;
}

// launches the consume kernel over all rows, then re-runs it over deferred rows after growing the
// table until every row found a slot
int run_consume(b200_groupby* gb, GbArgs& a) {
  b200_ctx* ctx = gb->ctx;
  const int64_t n = a.n;
  if (n == 0) return B200_OK;
  const size_t words = size_t((n + 31) / 32);
  const size_t claimed_bytes = a.mode == kModeInsert ? b200_align_up(size_t(n) * 4, 256) : 0;
  const size_t zf_bytes = size_t(ctx->num_sms) * 2 * kFastGroups * sizeof(uint64_t);
  const size_t spill_slots = std::max<size_t>(size_t(ctx->num_sms) * 2 * kFastGroups, size_t(ctx->num_sms) * (160 * 1024 / a.s.lstride));
  const size_t spill_bytes = a.mode == kModeFused ? spill_slots * size_t(a.s.stride) : 0;
  B200_RETURN_NOT_OK(gb_reserve_tmp(gb, 2 * b200_align_up(words * 4, 256) + claimed_bytes + zf_bytes + spill_bytes + 512));
  uint32_t* defer_a = reinterpret_cast<uint32_t*>(gb->tmp);
  uint32_t* defer_b = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(gb->tmp) + b200_align_up(words * 4, 256));
  if (a.mode == kModeInsert) a.claimed = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(gb->tmp) + 2 * b200_align_up(words * 4, 256));
  a.counters = gb->counters;
  a.defer_in = nullptr;
  a.defer_out = defer_a;
  B200_CUDA(ctx, cudaMemsetAsync(defer_a, 0, words * 4, ctx->stream));
  B200_CUDA(ctx, cudaMemsetAsync(gb->counters + 1, 0, 2 * sizeof(uint32_t), ctx->stream));
  a.spill = reinterpret_cast<char*>(gb->tmp) + 2 * b200_align_up(words * 4, 256) + claimed_bytes + 256 + zf_bytes;
  B200_CUDA(ctx, cudaMemsetAsync(gb->counters + 5, 0, sizeof(uint32_t), ctx->stream));

  // path choice: the shared-memory table pays off while the groups fit in it
  const int sm_slots_budget = 160 * 1024 / a.s.lstride;
  int local_slots = 1;
  while (local_slots * 2 <= sm_slots_budget) local_slots *= 2;
  const uint64_t likely_groups = std::max<uint64_t>(gb->groups, gb->expected);
  // (the grouper's insert/lookup passes stay on the global table: a growth here would move slots
  // whose positional ids are live)
  const bool low_card = a.mode == kModeFused && likely_groups * 2 <= uint64_t(local_slots) && local_slots >= 64 && !a.row_ids;
  // one value column (or none) and a cardinality hint (or history) that fits the shared table: the fast kernel.
  // With no hint at all the fast kernel is tried too: a CTA whose table fills up bypasses it.
  const bool fast = a.mode == kModeFused && a.n_cols == 1 && likely_groups <= uint64_t(kFastMaxGroupsHint) && !gb->fast_disabled && !a.row_ids;
  // (the partitioned path decides itself whether it needs the table)
  const bool try_hp = hp_applicable(gb, a, likely_groups);
  if (gb->pending || !try_hp) B200_RETURN_NOT_OK(materialize(gb));
  for (int round = 0; round < 8; ++round) {
    a.t = gb->table;
    if (round == 0 && try_hp) {
      bool handled = false;
      uint32_t hp_deferred = 0;
      B200_RETURN_NOT_OK(run_hp_consume(gb, a, likely_groups, &handled, &hp_deferred));
      if (!handled) B200_RETURN_NOT_OK(materialize(gb));
      if (handled) {
        if (hp_deferred == 0) return B200_OK;
        if (uint64_t(gb->groups) + hp_deferred > gb->table.cap / 2) B200_RETURN_NOT_OK(grow_table(gb, 2 * (uint64_t(gb->groups) + hp_deferred)));
        a.defer_in = defer_a;
        a.defer_out = defer_b;
        B200_CUDA(ctx, cudaMemsetAsync(defer_b, 0, words * 4, ctx->stream));
        B200_CUDA(ctx, cudaMemsetAsync(gb->counters + 1, 0, sizeof(uint32_t), ctx->stream));
        continue;
      }
      a.t = gb->table;
    }
    bool tiny = false;
    if (fast && round == 0 && likely_groups > 0 && likely_groups <= uint64_t(kTinyGroups)) {
      tiny = true;
      for (int j = 0; j < a.s.n_aggs; ++j)
        if (a.s.aggs[j].col == 0 && a.s.aggs[j].kind == B200_HASH_MIN_MAX) tiny = false;
    }
    if (tiny) {
      // <= 8 groups, sums / means / counts: thread-private accumulators (gb_tiny_kernel); rows it defers
      // (special keys, a 9th group) are picked up by the generic kernel below
      const int grid = int(std::min<int64_t>(int64_t(ctx->num_sms) * 4, (n + 4095) / 4096));
      int64_t rpb = (n + grid - 1) / grid;
      rpb = (rpb + 31) / 32 * 32;
      B200_RETURN_NOT_OK(grow_table(gb, 2 * std::max<uint64_t>(likely_groups, 32 * 1024)));
      a.t = gb->table;
      a.rows_per_block = int(rpb);
      if (b200_ctx_once(ctx, 4)) {
        B200_CUDA(ctx, cudaFuncSetAttribute(gb_tiny_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(TinySmem))));
        B200_CUDA(ctx, cudaFuncSetAttribute(gb_tiny8_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(TinySmem))));
        B200_CUDA(ctx, cudaFuncSetAttribute(gb_tiny8_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(TinySmem))));
        B200_CUDA(ctx, cudaFuncSetAttribute(gb_tiny8_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(TinySmem))));
        B200_CUDA(ctx, cudaFuncSetAttribute(gb_tiny8_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(TinySmem))));
      }
      // byte-wide keys without nulls and a 64-bit value column (the Q1 shape): register accumulators
      bool lean = a.s.n_keys <= 3 && !B200_KNOB("B200_GB_NO_TINY8");
      for (int k = 0; k < a.s.n_keys; ++k)
        if (a.keys[k].nbits != 8 || a.keys[k].vv.words != nullptr) lean = false;
      const int vt = a.cols[0].type;
      if (vt != B200_DOUBLE && vt != B200_INT64 && vt != B200_UINT64) lean = false;
      if (lean && vt == B200_DOUBLE && a.select)
        B200_TIMED(ctx, "groupby_consume", (gb_tiny8_kernel<true, true><<<grid, kTinyThreads, int(sizeof(TinySmem)), ctx->stream>>>(a)));
      else if (lean && vt == B200_DOUBLE)
        B200_TIMED(ctx, "groupby_consume", (gb_tiny8_kernel<true, false><<<grid, kTinyThreads, int(sizeof(TinySmem)), ctx->stream>>>(a)));
      else if (lean && a.select)
        B200_TIMED(ctx, "groupby_consume", (gb_tiny8_kernel<false, true><<<grid, kTinyThreads, int(sizeof(TinySmem)), ctx->stream>>>(a)));
      else if (lean)
        B200_TIMED(ctx, "groupby_consume", (gb_tiny8_kernel<false, false><<<grid, kTinyThreads, int(sizeof(TinySmem)), ctx->stream>>>(a)));
      else
        B200_TIMED(ctx, "groupby_consume", (gb_tiny_kernel<<<grid, kTinyThreads, int(sizeof(TinySmem)), ctx->stream>>>(a)));
      uint32_t c[4];
      B200_RETURN_NOT_OK(read_counters(gb, c));
      const uint32_t tiny_deferred = c[1];
      B200_RETURN_NOT_OK(import_spilled(gb, a.spill));
      if (uint64_t(gb->groups) > uint64_t(kFastMaxGroupsHint)) gb->fast_disabled = true;
      if (tiny_deferred == 0) {
        if (uint64_t(gb->groups) * 2 > gb->table.cap) B200_RETURN_NOT_OK(grow_table(gb, uint64_t(gb->groups) * 4));
        return B200_OK;
      }
      if (uint64_t(gb->groups) + tiny_deferred > gb->table.cap / 2) B200_RETURN_NOT_OK(grow_table(gb, 2 * (uint64_t(gb->groups) + tiny_deferred)));
      a.defer_in = defer_a;
      a.defer_out = defer_b;
      B200_CUDA(ctx, cudaMemsetAsync(defer_b, 0, words * 4, ctx->stream));
      B200_CUDA(ctx, cudaMemsetAsync(gb->counters + 1, 0, sizeof(uint32_t), ctx->stream));
      continue;
    }
    if (fast && !a.select && round == 0) {
      // full 2048-row tiles go through the fast kernel; the ragged tail and every row it defers
      // (special keys, keys that found no room in a CTA's table) are taken by the generic kernel
      // v2 (binned, one 1024-thread CTA per SM) unless B200_GB_V1 / B200_GB_DIRECT ask for the earlier kernels
      // The binned kernels want many groups: a group may hold kBinCap values per 8192-row tile before rows
      // take the slow overflow path, so below ~512 expected groups (>= 16 rows per group per tile) the
      // counting-sort kernel, whose bins are sized exactly by a scan, is the right one.
      static const bool binned_ok = !B200_KNOB("B200_GB_V1") && !B200_KNOB("B200_GB_DIRECT");
      const bool binned = binned_ok && likely_groups * kBinCap >= uint64_t(kBinTile);
      const int64_t tile_rows = binned ? kBinTile : kFastTile;
      const int64_t n_full = n / tile_rows * tile_rows;
      if (n_full < n) B200_CUDA(ctx, cudaMemsetAsync(defer_a + n_full / 32, 0xFF, (words - size_t(n_full / 32)) * 4, ctx->stream));
      uint32_t fast_deferred = 0;
      if (n_full > 0) {
        const int max_blocks = binned ? ctx->num_sms : ctx->num_sms * 2;
        int64_t rpb = (n_full + max_blocks - 1) / max_blocks;
        rpb = (rpb + tile_rows - 1) / tile_rows * tile_rows;
        const int grid = int((n_full + rpb - 1) / rpb);
        B200_RETURN_NOT_OK(grow_table(gb, 2 * std::max<uint64_t>(likely_groups, 32 * 1024)));
        a.t = gb->table;

        a.local_slots = kFastSlots32;
        static const int dbg = B200_KNOB("B200_GB_DBG") ? atoi(B200_KNOB("B200_GB_DBG")) : 0;
        a.dbg = dbg;
        a.rows_per_block = int(rpb);
        a.n = n_full;
        a.zf_scratch = reinterpret_cast<uint64_t*>(reinterpret_cast<char*>(gb->tmp) + 2 * b200_align_up(words * 4, 256) + claimed_bytes + 256);
        B200_CUDA(ctx, cudaMemsetAsync(a.zf_scratch, 0xFF, size_t(grid) * kFastGroups * sizeof(uint64_t), ctx->stream));
        const int smem = binned ? int(sizeof(BinSmem)) : int(sizeof(FastSmem));
        const bool k32 = a.s.n_keys == 1 && a.keys[0].nbits == 32 && !a.keys[0].vv.words && (reinterpret_cast<uintptr_t>(a.keys[0].data) & 15) == 0;
        const bool k64 = a.s.n_keys == 1 && a.keys[0].nbits == 64 && !a.keys[0].vv.words && (reinterpret_cast<uintptr_t>(a.keys[0].data) & 15) == 0;
        const bool vfast = a.n_cols == 1 && b200_type_bits(a.cols[0].type) == 64 && (reinterpret_cast<uintptr_t>(a.cols[0].data) & (B200_BIN_LD == 2 ? 31 : 15)) == 0;
        if (b200_ctx_once(ctx, 5)) {
          const int mx = int(sizeof(FastSmem));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_fast_kernel<0, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_fast_kernel<0, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_fast_kernel<0, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_fast_kernel<0, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_fast_kernel<1, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_fast_kernel<1, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_fast_kernel<1, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_fast_kernel<1, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_fast_kernel<2, false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_fast_kernel<2, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_fast_kernel<2, true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_fast_kernel<2, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, mx));
          const int bx = int(sizeof(BinSmem));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_bin_kernel<0, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_bin_kernel<0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_bin_kernel<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_bin_kernel<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_bin_kernel<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, bx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_bin_kernel<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, bx));
          B200_CUDA(ctx, cudaFuncSetAttribute(gb_bin32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bx));
        }
        static const bool direct = B200_KNOB("B200_GB_DIRECT") != nullptr;
#define B200_FAST_LAUNCH(KM, VF) \
  do {                                                                                                              \
    if (binned) B200_TIMED(ctx, "groupby_consume", (gb_bin_kernel<KM, VF><<<grid, kBinThreads, smem, ctx->stream>>>(a)));          \
    else if (direct) B200_TIMED(ctx, "groupby_consume", (gb_fast_kernel<KM, VF, true><<<grid, kFastThreads, smem, ctx->stream>>>(a)));  \
    else B200_TIMED(ctx, "groupby_consume", (gb_fast_kernel<KM, VF, false><<<grid, kFastThreads, smem, ctx->stream>>>(a)));        \
  } while (0)
        static const bool pipelined = B200_KNOB("B200_GB_V2") == nullptr;
        // (value validity at a bit offset that is a multiple of 4: the four bits of a thread's rows then never leave
        // one 16-bit word, which is all gb_bin32_kernel keeps of them; other offsets take gb_bin_kernel)
        if (binned && pipelined && k32 && vfast && a.keys[0].nbits == 32 && (a.cols[0].vv.bit_offset & 3) == 0) {
          B200_TIMED(ctx, "groupby_consume", (gb_bin32_kernel<<<grid, kBinThreads, smem, ctx->stream>>>(a)));
        } else if (k32) { if (vfast) B200_FAST_LAUNCH(1, true); else B200_FAST_LAUNCH(1, false); }
        else if (k64) { if (vfast) B200_FAST_LAUNCH(2, true); else B200_FAST_LAUNCH(2, false); }
        else { if (vfast) B200_FAST_LAUNCH(0, true); else B200_FAST_LAUNCH(0, false); }
#undef B200_FAST_LAUNCH
        a.n = n;
        uint32_t c[4];
        B200_RETURN_NOT_OK(read_counters(gb, c));
        fast_deferred = c[1];
        B200_RETURN_NOT_OK(import_spilled(gb, a.spill));
        if (uint64_t(gb->groups) > uint64_t(kFastMaxGroupsHint)) gb->fast_disabled = true;
      }
      if (fast_deferred == 0 && n_full == n) {
        if (uint64_t(gb->groups) * 2 > gb->table.cap) B200_RETURN_NOT_OK(grow_table(gb, uint64_t(gb->groups) * 4));
        return B200_OK;
      }
      // the generic kernel below picks up exactly the flagged rows
      if (uint64_t(gb->groups) + fast_deferred + uint64_t(n - n_full) > gb->table.cap / 2)
        B200_RETURN_NOT_OK(grow_table(gb, 2 * (uint64_t(gb->groups) + fast_deferred + uint64_t(n - n_full))));
      a.defer_in = defer_a;
      a.defer_out = defer_b;
      B200_CUDA(ctx, cudaMemsetAsync(defer_b, 0, words * 4, ctx->stream));
      B200_CUDA(ctx, cudaMemsetAsync(gb->counters + 1, 0, sizeof(uint32_t), ctx->stream));
      continue;
    } else if (low_card && round == 0) {
      const int blocks = ctx->num_sms;
      // sized from the hint; a flush that finds no room spills (see import_spilled)
      B200_RETURN_NOT_OK(grow_table(gb, 2 * std::max<uint64_t>(likely_groups, 32 * 1024)));
      a.t = gb->table;
      a.local_slots = local_slots;
      a.rows_per_block = int((n + blocks - 1) / blocks);
      a.rows_per_block = (a.rows_per_block + 1023) / 1024 * 1024;
      const int smem = local_slots * a.s.lstride;
      if (b200_ctx_once(ctx, 2)) {
        B200_CUDA(ctx, cudaFuncSetAttribute(gb_consume_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      }
      B200_TIMED(ctx, "groupby_consume", (gb_consume_kernel<true><<<blocks, 1024, smem, ctx->stream>>>(a)));
    } else {
      a.local_slots = 1;
      a.rows_per_block = 4096;
      const int64_t blocks = (n + a.rows_per_block - 1) / a.rows_per_block;
      B200_TIMED(ctx, "groupby_consume", (gb_consume_kernel<false><<<(unsigned)blocks, 256, 0, ctx->stream>>>(a)));
    }
    uint32_t c[4];
    B200_RETURN_NOT_OK(read_counters(gb, c));
    const uint32_t deferred = c[1];
    B200_RETURN_NOT_OK(import_spilled(gb, a.spill));
    // cardinality beyond the shared table: later batches go straight to the global table
    if (uint64_t(gb->groups) > uint64_t(kFastMaxGroupsHint)) gb->fast_disabled = true;
    if (deferred == 0) {
      // keep the load factor <= 1/2 for the next batch
      if (uint64_t(gb->groups) * 2 > gb->table.cap) B200_RETURN_NOT_OK(grow_table(gb, uint64_t(gb->groups) * 4));
      return B200_OK;
    }
    if (a.mode == kModeInsert && gb->slot_ids) {
      // a rehash invalidates positional ids and the claimed list: the grouper restarts this batch
      return b200_fail(ctx, B200_CAPACITY_ERROR, "grouper: table overflow");
    }
    B200_RETURN_NOT_OK(grow_table(gb, std::max<uint64_t>(gb->table.cap * 4, 4 * (uint64_t(gb->groups) + deferred))));
    a.defer_in = a.defer_out;
    a.defer_out = (a.defer_in == defer_a) ? defer_b : defer_a;
    B200_CUDA(ctx, cudaMemsetAsync(a.defer_out, 0, words * 4, ctx->stream));
    B200_CUDA(ctx, cudaMemsetAsync(gb->counters + 1, 0, sizeof(uint32_t), ctx->stream));
  }
  return b200_fail(ctx, B200_CAPACITY_ERROR, "groupby: rows still deferred after 8 table growths");
}

}  // namespace

extern "C" {

int b200_groupby_create(b200_ctx* ctx, int n_keys, const int32_t* key_types, int n_aggs, const b200_agg_spec* aggs,
                        int64_t expected_groups, b200_groupby** out) {
  B200_ENTER(ctx);
  if (!ctx || !out) return B200_INVALID;
  *out = nullptr;
  if (n_keys < 1 || n_keys > kWideMaxKeys)
    return b200_fail(ctx, B200_NOT_IMPLEMENTED, "groupby: %d key columns (1..%d supported)", n_keys, kWideMaxKeys);
  if (n_aggs < 0 || n_aggs > kMaxAggs) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "groupby: %d aggregates (max %d)", n_aggs, kMaxAggs);
  GbSchema s;
  memset(&s, 0, sizeof(s));
  s.n_aggs = n_aggs;
  int total = 0;
  bool any_var = false;   // a variable-length key column: always through the key -> id table
  for (int k = 0; k < n_keys; ++k) {
    // GrouperImpl::Make: "Keys of type ..." NotImplemented for unsupported types (hash_aggregate.cc:352)
    if (key_types[k] == B200_BINARY) {
      any_var = true;
      continue;
    }
    if (b200_type_bits(key_types[k]) == 0) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "Keys of type %d", key_types[k]);
    total += b200_type_bits(key_types[k]);
  }
  // Keys that fit one 64-bit register (values + one null flag per column, or a single 64-bit column) are packed;
  // anything wider goes through the key -> id table of groupby_wide.cuh and is grouped by a uint32 id column.
  const int wide_n = n_keys;
  const bool wide = any_var || n_keys > kMaxKeys || !(total + n_keys <= 64 || (n_keys == 1 && total == 64));
  int32_t wide_types[kWideMaxKeys];
  if (wide) {
    for (int k = 0; k < n_keys; ++k) wide_types[k] = key_types[k];
    static const int32_t id_type = B200_UINT32;
    key_types = &id_type;
    n_keys = 1;
    total = 32;
  }
  s.n_keys = n_keys;
  total = 0;
  for (int k = 0; k < n_keys; ++k) {
    const int bits = b200_type_bits(key_types[k]);
    s.key_type[k] = key_types[k];
    s.key_bits[k] = bits;
    s.key_shift[k] = total;
    total += bits;
  }
  s.null_shift = (total + n_keys <= 64) ? total : -1;
  int off = 16;
  for (int j = 0; j < n_aggs; ++j) {
    const int kind = aggs[j].kind, t = aggs[j].arg_type;
    if (kind < B200_HASH_COUNT || kind > B200_HASH_MEAN) return b200_fail(ctx, B200_KEY_ERROR, "No function registered with kind %d", kind);
    if (b200_type_bits(t) == 0) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "groupby: argument type %d", t);
    // GroupedSumImpl / GroupedMinMaxImpl: "Summing data of type" / "Grouped MinMax data of type" NotImplemented
    // (hash_aggregate.cc:550-556, 684-694)
    if (kind != B200_HASH_COUNT && t == B200_BOOL)
      return b200_fail(ctx, B200_NOT_IMPLEMENTED, "%s data of type bool", kind == B200_HASH_MIN_MAX ? "Grouped MinMax" : "Summing");
    s.aggs[j].kind = kind;
    s.aggs[j].dom = domain_of(t);
    s.aggs[j].option = aggs[j].option;
    s.aggs[j].off = off;
    s.aggs[j].col = j;
    s.aggs[j].arg_type = t;
    off += state_bytes(kind);
  }
  s.lstride = off;
  s.stride = int(b200_align_up(size_t(off) + 8, 16));  // + one padding word (export tag)
  b200_groupby* gb = new b200_groupby();
  gb->ctx = ctx;
  gb->schema = s;
  gb->specs.assign(aggs, aggs + n_aggs);
  gb->expected = expected_groups > 0 ? uint64_t(expected_groups) : 0;
  if (wide) {
    gb->wide = true;
    int pos = 0;
    gb->wn_keys = wide_n;
    for (int k = 0; k < wide_n; ++k) {
      gb->wtypes[k] = wide_types[k];
      gb->wbits[k] = b200_type_bits(wide_types[k]);
      gb->woff[k] = pos;
      gb->wvar_index[k] = -1;
      if (wide_types[k] == B200_BINARY) {   // [null byte][4-byte length]
        gb->wvar_index[k] = gb->wn_var;
        gb->wvar_col[gb->wn_var++] = k;
        pos += 5;
      } else {
        pos += 1 + (gb->wbits[k] == 1 ? 1 : gb->wbits[k] / 8);
      }
    }
    gb->wwords = (pos + 7) / 8;
  }
  if (b200_dev_alloc(ctx, reinterpret_cast<void**>(&gb->counters), 8 * sizeof(uint32_t)) != B200_OK) {
    delete gb;
    return B200_OUT_OF_MEMORY;
  }
  cudaMemsetAsync(gb->counters, 0, 8 * sizeof(uint32_t), ctx->stream);
  // with a hint in the partitioned path's range the table is allocated when first needed (b200_groupby::pending)
  if (int64_t(gb->expected) < gb->hp_min_groups) {
    int st = alloc_table(gb, initial_cap(gb), &gb->table);
    if (st != B200_OK) {
      b200_dev_free(ctx, gb->counters);
      delete gb;
      return st;
    }
  }
  *out = gb;
  return B200_OK;
}

int b200_groupby_destroy(b200_groupby* gb) {
  B200_ENTER(gb ? gb->ctx : nullptr);
  if (!gb) return B200_OK;
  cudaStreamSynchronize(gb->ctx->stream);
  b200_dev_free(gb->ctx, gb->table.slots);
  b200_dev_free(gb->ctx, gb->counters);
  b200_dev_free(gb->ctx, gb->slot_ids);
  b200_dev_free(gb->ctx, gb->dense);
  b200_dev_free(gb->ctx, gb->tmp);
  b200_dev_free(gb->ctx, gb->hp_tmp);
  b200_dev_free(gb->ctx, gb->wslots);
  b200_dev_free(gb->ctx, gb->wstore);
  b200_dev_free(gb->ctx, gb->wcounters);
  b200_dev_free(gb->ctx, gb->wtmp);
  b200_dev_free(gb->ctx, gb->wheap);
  b200_dev_free(gb->ctx, gb->wstore_var);
  b200_dev_free(gb->ctx, gb->wvar_bytes);
  delete gb;
  return B200_OK;
}


// ---- wide keys: key columns -> one uint32 column of key ids (groupby_wide.cuh) ------------------------------------
namespace {
__global__ void wide_fill_kernel(uint64_t* p, uint64_t n) {
  const uint64_t i = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) p[i] = kWideEmpty;
}

int wide_translate(b200_groupby* gb, const b200_col* keys, const uint32_t* select, b200_col* out) {
  b200_ctx* ctx = gb->ctx;
  const int64_t n = keys[0].length;
  WideArgs a;
  memset(&a, 0, sizeof(a));
  a.n_keys = gb->wn_keys;
  a.words = gb->wwords;
  a.n = n;
  a.select = select;
  for (int k = 0; k < gb->wn_keys; ++k) {
    const b200_col& c = keys[k];
    if (c.type != gb->wtypes[k]) return b200_fail(ctx, B200_TYPE_ERROR, "groupby: key %d has type %d, expected %d", k, c.type, gb->wtypes[k]);
    if (c.length != n) return b200_fail(ctx, B200_INVALID, "Array arguments must all be the same length");
    if (n > 0 && !c.data) return b200_fail(ctx, B200_INVALID, "groupby: key %d has no data buffer", k);
    WideKeyCol& w = a.cols[k];
    const int bits = gb->wbits[k];
    w.nbits = bits;
    w.byte_off = gb->woff[k];
    w.var_index = gb->wvar_index[k];
    if (c.type == B200_BINARY) {
      // (a column of empty strings may come without a value buffer: any non-null pointer will do, nothing is read)
      w.bytes = reinterpret_cast<const uint8_t*>(c.extra ? c.extra : c.data);
      w.offsets = reinterpret_cast<const int32_t*>(c.data) + c.offset;
    }
    w.data = bits >= 8 ? reinterpret_cast<const char*>(c.data) + c.offset * (bits / 8) : nullptr;
    w.bits = make_bitmap_ref(bits == 1 ? c.data : nullptr, c.offset);
    const bool has_nulls = c.validity != nullptr && c.null_count != 0;
    w.vv = has_nulls ? make_bitmap_ref(c.validity, c.offset) : BitmapRef{nullptr, 0};
  }
  memset(out, 0, sizeof(*out));
  out->type = B200_UINT32;
  out->length = n;
  if (n == 0) return B200_OK;
  if (n >= (int64_t(1) << 32)) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "groupby: more than 2^32-1 rows per call with wide keys");
  if (!gb->wcounters) {
    B200_RETURN_NOT_OK(b200_dev_alloc(ctx, reinterpret_cast<void**>(&gb->wcounters), 64));
    B200_CUDA(ctx, cudaMemsetAsync(gb->wcounters, 0, 64, ctx->stream));
  }
  // table at load <= 1/2 even if every row brings a new key; a grown table gets the stored keys back
  uint64_t want = 1024;
  while (want < 2 * (uint64_t(gb->wcount) + uint64_t(n))) want <<= 1;
  if (want > gb->wcap) {
    if (gb->wslots) b200_dev_free(ctx, gb->wslots);
    gb->wslots = nullptr;
    B200_RETURN_NOT_OK(b200_dev_alloc(ctx, reinterpret_cast<void**>(&gb->wslots), want * 8));
    gb->wcap = want;
    wide_fill_kernel<<<(unsigned)((want + 255) / 256), 256, 0, ctx->stream>>>(gb->wslots, want);
    B200_LAUNCHED(ctx);
    if (gb->wcount) {
      WideVarInfo vi;
      memset(&vi, 0, sizeof(vi));
      vi.n_var = gb->wn_var;
      for (int v = 0; v < gb->wn_var; ++v) vi.byte_off[v] = gb->woff[gb->wvar_col[v]];
      vi.store_var = gb->wstore_var;
      vi.heap = gb->wheap;
      wide_rehash_kernel<<<(gb->wcount + 255) / 256, 256, 0, ctx->stream>>>(gb->wstore, gb->wwords, gb->wcount, gb->wslots, gb->wcap, gb->wcounters, vi);
      B200_LAUNCHED(ctx);
    }
  }
  const size_t a4 = b200_align_up(size_t(n) * 4, 256);
  const size_t tmp_need = 3 * a4 + (gb->wn_var ? a4 + 256 : 0);
  if (tmp_need > gb->wtmp_bytes) {
    if (gb->wtmp) b200_dev_free(ctx, gb->wtmp);
    gb->wtmp = nullptr;
    gb->wtmp_bytes = 0;
    B200_RETURN_NOT_OK(b200_dev_alloc(ctx, &gb->wtmp, tmp_need));
    gb->wtmp_bytes = tmp_need;
  }
  if (gb->wn_var && !gb->wvar_bytes) {
    B200_RETURN_NOT_OK(b200_dev_alloc(ctx, reinterpret_cast<void**>(&gb->wvar_bytes), kWideMaxKeys * sizeof(unsigned long long)));
    B200_CUDA(ctx, cudaMemsetAsync(gb->wvar_bytes, 0, kWideMaxKeys * sizeof(unsigned long long), ctx->stream));
  }
  a.var_bytes = gb->wvar_bytes;
  a.new_len = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(gb->wtmp) + 3 * a4);
  a.n_var = gb->wn_var;
  for (int v = 0; v < gb->wn_var; ++v) a.var_col[v] = gb->wvar_col[v];
  a.store_var = gb->wstore_var;
  a.heap = gb->wheap;
  a.var_bytes = gb->wvar_bytes;
  a.slots = gb->wslots;
  a.cap = gb->wcap;
  a.store = gb->wstore;
  a.counters = gb->wcounters;
  a.new_slots = reinterpret_cast<uint32_t*>(gb->wtmp);
  a.row_slot = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(gb->wtmp) + a4);
  a.ids = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(gb->wtmp) + 2 * a4);
  B200_CUDA(ctx, cudaMemsetAsync(gb->wcounters + 1, 0, sizeof(uint32_t), ctx->stream));
  const unsigned grid = unsigned((n + 255) / 256);
  B200_TIMED(ctx, "groupby_wide_keys", (wide_insert_kernel<<<grid, 256, 0, ctx->stream>>>(a)));
  B200_CUDA(ctx, cudaMemcpyAsync(ctx->host_words + 40, gb->wcounters, 4 * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
  B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  const uint32_t* hc = reinterpret_cast<const uint32_t*>(ctx->host_words + 40);
  if (hc[2] != 0) return b200_fail(ctx, B200_UNKNOWN_ERROR, "groupby: wide-key table error %u", hc[2]);
  const uint32_t n_new = hc[1];
  if (n_new) {
    if (uint64_t(gb->wcount) + n_new > gb->wstore_cap) {
      uint64_t cap = std::max<uint64_t>(1024, gb->wstore_cap);
      while (cap < uint64_t(gb->wcount) + n_new) cap <<= 1;
      if (cap > 0xFFFFFFF0ull) return b200_fail(ctx, B200_CAPACITY_ERROR, "groupby: more than 2^32 distinct wide keys");
      uint64_t* ns = nullptr;
      B200_RETURN_NOT_OK(b200_dev_alloc(ctx, reinterpret_cast<void**>(&ns), cap * gb->wwords * 8));
      if (gb->wcount) B200_CUDA(ctx, cudaMemcpyAsync(ns, gb->wstore, size_t(gb->wcount) * gb->wwords * 8, cudaMemcpyDeviceToDevice, ctx->stream));
      if (gb->wstore) {
        B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        b200_dev_free(ctx, gb->wstore);
      }
      gb->wstore = ns;
      a.store = ns;
      if (gb->wn_var) {
        uint64_t* nv = nullptr;
        B200_RETURN_NOT_OK(b200_dev_alloc(ctx, reinterpret_cast<void**>(&nv), cap * gb->wn_var * 8));
        if (gb->wcount) B200_CUDA(ctx, cudaMemcpyAsync(nv, gb->wstore_var, size_t(gb->wcount) * gb->wn_var * 8, cudaMemcpyDeviceToDevice, ctx->stream));
        if (gb->wstore_var) {
          B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
          b200_dev_free(ctx, gb->wstore_var);
        }
        gb->wstore_var = nv;
        a.store_var = nv;
      }
      gb->wstore_cap = uint32_t(cap);
    }
    if (gb->wn_var) {
      // bytes the new keys add to the heap: lengths -> scan -> total on the host -> room in the heap
      wide_new_len_kernel<<<(n_new + 255) / 256, 256, 0, ctx->stream>>>(a, n_new);
      B200_LAUNCHED(ctx);
      gb_hp_scan_kernel<<<1, 1024, 0, ctx->stream>>>(a.new_len, n_new, nullptr, a.new_len);
      B200_LAUNCHED(ctx);
      B200_CUDA(ctx, cudaMemcpyAsync(ctx->host_words + 44, a.new_len + n_new, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
      B200_CUDA(ctx, cudaMemcpyAsync(ctx->host_words + 46, gb->wvar_bytes, kWideMaxKeys * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
      B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
      const uint64_t add = *reinterpret_cast<const uint32_t*>(ctx->host_words + 44);
      for (int v = 0; v < gb->wn_var; ++v) gb->wvar_bytes_host[v] = uint64_t(ctx->host_words[46 + v]);
      // (a batch's value buffers are addressed by int32 offsets: its new keys cannot add 4 GB, the scan cannot wrap)
      if (gb->wheap_used + add > gb->wheap_cap) {
        uint64_t cap = std::max<uint64_t>(1 << 20, gb->wheap_cap);
        while (cap < gb->wheap_used + add) cap <<= 1;
        uint8_t* nh = nullptr;
        B200_RETURN_NOT_OK(b200_dev_alloc(ctx, reinterpret_cast<void**>(&nh), cap));
        if (gb->wheap_used) B200_CUDA(ctx, cudaMemcpyAsync(nh, gb->wheap, gb->wheap_used, cudaMemcpyDeviceToDevice, ctx->stream));
        if (gb->wheap) {
          B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
          b200_dev_free(ctx, gb->wheap);
        }
        gb->wheap = nh;
        gb->wheap_cap = cap;
      }
      a.heap = gb->wheap;
      wide_assign_kernel<<<(n_new + 255) / 256, 256, 0, ctx->stream>>>(a, n_new, gb->wcount, gb->wheap_used);
      B200_LAUNCHED(ctx);
      gb->wheap_used += add;
    } else {
      wide_assign_kernel<<<(n_new + 255) / 256, 256, 0, ctx->stream>>>(a, n_new, gb->wcount, 0);
      B200_LAUNCHED(ctx);
    }
    gb->wcount += n_new;
  }
  B200_TIMED(ctx, "groupby_wide_keys", (wide_ids_kernel<<<grid, 256, 0, ctx->stream>>>(a)));
  out->data = a.ids;
  out->null_count = 0;
  return B200_OK;
}
}  // namespace

#define B200_NARROW_KEYS(gb, keys, select)                                   \
  b200_col _narrow_key;                                                      \
  if ((gb) && (gb)->wide && (keys)) {                                        \
    B200_RETURN_NOT_OK(wide_translate((gb), (keys), (select), &_narrow_key)); \
    (keys) = &_narrow_key;                                                   \
  }

int b200_groupby_key_bytes(b200_groupby* gb, int key_index, int64_t* out_bytes) {
  if (!gb || !out_bytes) return B200_INVALID;
  *out_bytes = 0;
  if (!gb->wide) return key_index >= 0 && key_index < gb->schema.n_keys ? B200_OK : b200_fail(gb->ctx, B200_INVALID, "groupby_key_bytes: no key %d", key_index);
  if (key_index < 0 || key_index >= gb->wn_keys) return b200_fail(gb->ctx, B200_INVALID, "groupby_key_bytes: no key %d", key_index);
  if (gb->wvar_index[key_index] >= 0) {
    const uint64_t b = gb->wvar_bytes_host[gb->wvar_index[key_index]];
    // BinaryType's offsets are int32 (VarLengthKeyEncoder::Decode sums into Offset, hash_aggregate.cc:282-285)
    if (b > 0x7FFFFFFFull) return b200_fail(gb->ctx, B200_CAPACITY_ERROR, "groupby: %llu bytes of distinct keys in column %d exceed int32 offsets", (unsigned long long)b, key_index);
    *out_bytes = int64_t(b);
  }
  return B200_OK;
}

int b200_groupby_set_option(b200_groupby* gb, const char* name, int64_t value) {
  B200_ENTER(gb ? gb->ctx : nullptr);
  if (!gb || !name) return B200_INVALID;
  if (!strcmp(name, "hp_min_rows")) gb->hp_min_rows = value;
  else if (!strcmp(name, "hp_min_groups")) gb->hp_min_groups = value;
  else if (!strcmp(name, "hp_keys_per_part")) gb->hp_keys_per_part = std::max<int64_t>(64, std::min<int64_t>(value, kHpSlots / 2));
  else return b200_fail(gb->ctx, B200_KEY_ERROR, "groupby: no option named %s", name);
  return B200_OK;
}

int b200_groupby_stat(b200_groupby* gb, const char* name, int64_t* out) {
  if (!gb || !name || !out) return B200_INVALID;
  if (!strcmp(name, "partitioned_calls")) *out = gb->hp_calls;
  else if (!strcmp(name, "partitioned_alloc_failures")) *out = gb->hp_alloc_failures;
  else if (!strcmp(name, "partitioned_span_refusals")) *out = gb->hp_span_refusals;
  else return b200_fail(gb->ctx, B200_KEY_ERROR, "groupby: no statistic named %s", name);
  return B200_OK;
}

static int groupby_consume_impl(b200_groupby* gb, const b200_col* keys, const b200_col* args, uint64_t row_base,
                                const uint64_t* row_ids, const uint32_t* select = nullptr);

int b200_groupby_consume(b200_groupby* gb, const b200_col* keys, const b200_col* args, uint64_t row_base) {
  B200_ENTER(gb ? gb->ctx : nullptr);
  return groupby_consume_impl(gb, keys, args, row_base, nullptr);
}

int b200_groupby_consume_selected(b200_groupby* gb, const b200_col* keys, const b200_col* args, uint64_t row_base,
                                  const void* selection) {
  B200_ENTER(gb ? gb->ctx : nullptr);
  if (gb && !selection && keys && keys[0].length > 0) return b200_fail(gb->ctx, B200_INVALID, "groupby_consume_selected: null selection");
  if (gb && (reinterpret_cast<uintptr_t>(selection) & 3)) return b200_fail(gb->ctx, B200_INVALID, "groupby_consume_selected: selection must be 4-byte aligned");
  return groupby_consume_impl(gb, keys, args, row_base, nullptr, reinterpret_cast<const uint32_t*>(selection));
}

int b200_groupby_consume_rows(b200_groupby* gb, const b200_col* keys, const b200_col* args, const uint64_t* row_ids) {
  B200_ENTER(gb ? gb->ctx : nullptr);
  if (gb && !row_ids && keys && keys[0].length > 0) return b200_fail(gb->ctx, B200_INVALID, "groupby_consume_rows: null row_ids");
  return groupby_consume_impl(gb, keys, args, 0, row_ids);
}

static int groupby_consume_impl(b200_groupby* gb, const b200_col* keys, const b200_col* args, uint64_t row_base,
                                const uint64_t* row_ids, const uint32_t* select) {
  if (!gb) return B200_INVALID;
  b200_ctx* ctx = gb->ctx;
  if (!keys || (gb->schema.n_aggs > 0 && !args)) return b200_fail(ctx, B200_INVALID, "groupby_consume: null argument");
  if (gb->dense_used) return b200_fail(ctx, B200_INVALID, "groupby: handle already used through b200_hash_agg_consume");
  gb->order_cache = nullptr;
  B200_NARROW_KEYS(gb, keys, select);
  GbArgs a;
  memset(&a, 0, sizeof(a));
  a.s = gb->schema;
  B200_RETURN_NOT_OK(fill_keys(gb, keys, &a));
  // de-duplicate argument columns: the same buffers feed several aggregates (C3, Q1)
  a.n_cols = 0;
  for (int j = 0; j < a.s.n_aggs; ++j) {
    const b200_col& c = args[j];
    if (c.type != a.s.aggs[j].arg_type) return b200_fail(ctx, B200_TYPE_ERROR, "groupby: argument %d has type %d, expected %d", j, c.type, a.s.aggs[j].arg_type);
    if (c.length != a.n) return b200_fail(ctx, B200_INVALID, "Array arguments must all be the same length");
    if (a.n > 0 && !c.data) return b200_fail(ctx, B200_INVALID, "groupby: argument %d has no data buffer", j);
    int found = -1;
    for (int p = 0; p < j; ++p)
      if (args[p].data == c.data && args[p].validity == c.validity && args[p].offset == c.offset && args[p].type == c.type &&
          ((args[p].null_count != 0) == (c.null_count != 0)))
        found = a.s.aggs[p].col;
    if (found < 0) {
      if (a.n_cols == kMaxCols) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "groupby: more than %d distinct argument columns", kMaxCols);
      fill_col(c, &a.cols[a.n_cols]);
      found = a.n_cols++;
    }
    a.s.aggs[j].col = found;
  }
  a.mode = kModeFused;
  a.row_base = row_base;
  a.row_ids = row_ids;
  a.select = select;
  gb->row_bound = row_ids ? ~uint64_t(0) : std::max(gb->row_bound, row_base + uint64_t(a.n));
  // Several distinct argument columns at low cardinality (the Q1 shape): aggregates over different columns
  // are independent, so the batch is consumed once per column with the binned single-column kernel — the
  // key columns are re-read (2 B/row for Q1) instead of funnelling every column's float sums through CAS
  // loops on a handful of shared-memory slots (measured 92 ms vs 3 x ~5 ms at 600 M rows, 4 groups).
  const uint64_t likely = std::max<uint64_t>(gb->groups, gb->expected);
  if (a.n_cols > 1 && !row_ids && likely > 0 && likely <= uint64_t(kFastMaxGroupsHint) && !gb->fast_disabled && a.n >= kBinTile) {
    for (int c = 0; c < a.n_cols; ++c) {
      GbArgs b = a;
      b.n_cols = 1;
      b.cols[0] = a.cols[c];
      for (int j = 0; j < b.s.n_aggs; ++j) b.s.aggs[j].col = a.s.aggs[j].col == c ? 0 : -1;
      B200_RETURN_NOT_OK(run_consume(gb, b));
    }
    return B200_OK;
  }
  return run_consume(gb, a);
}


// ---- fused compare -> filter -> derive -> aggregate ----------------------------------------------------
namespace {
bool same_col(const b200_col& x, const b200_col& y) {
  return x.data == y.data && x.validity == y.validity && x.offset == y.offset && x.type == y.type && x.length == y.length;
}
bool col_has_nulls(const b200_col& c) { return c.validity != nullptr && c.null_count != 0; }

// the one-pass kernel: *taken = false when the shape is not its own or a CTA met a 9th group (nothing was changed)
int try_fused8(b200_groupby* gb, const b200_col* keys, const b200_fused_arg* args, uint64_t row_base, const b200_predicate* pred,
               bool* taken) {
  b200_ctx* ctx = gb->ctx;
  *taken = false;
  const GbSchema& s = gb->schema;
  const uint64_t likely = std::max<uint64_t>(gb->groups, gb->expected);
  if (gb->wide || likely == 0 || likely > uint64_t(kTinyGroups) || s.n_keys > 3 || s.n_aggs == 0) return B200_OK;
  const int64_t n = keys[0].length;
  if (n < kBinTile) return B200_OK;
  for (int k = 0; k < s.n_keys; ++k)
    if (s.key_bits[k] != 8 || col_has_nulls(keys[k])) return B200_OK;
  if (pred) {
    if ((pred->column.type != B200_INT32 && pred->column.type != B200_UINT32) || col_has_nulls(pred->column)) return B200_OK;
  }
  FusedArgs f;
  memset(&f, 0, sizeof(f));
  b200_col cols[kFusedCols];
  auto col_index = [&](const b200_col& c) -> int {
    for (int i = 0; i < f.n_cols; ++i)
      if (same_col(cols[i], c)) return i;
    if (f.n_cols == kFusedCols) return -1;
    cols[f.n_cols] = c;
    f.cols[f.n_cols] = reinterpret_cast<const double*>(c.data) + c.offset;
    return f.n_cols++;
  };
  GbArgs a;
  memset(&a, 0, sizeof(a));
  a.s = s;
  B200_RETURN_NOT_OK(fill_keys(gb, keys, &a));
  // streams: the plain columns first (stream i reads column i), then the derived arguments
  for (int pass = 0; pass < 2; ++pass) {
    for (int j = 0; j < s.n_aggs; ++j) {
      const b200_fused_arg& g = args[j];
      if (g.kind != pass) continue;
      if (s.aggs[j].kind == B200_HASH_MIN_MAX) return B200_OK;
      if (g.a.type != B200_DOUBLE || col_has_nulls(g.a) || g.a.length != n) return B200_OK;
      if (g.kind == 1 && (g.b.type != B200_DOUBLE || col_has_nulls(g.b) || g.b.length != n)) return B200_OK;
      FusedStream st;
      memset(&st, 0, sizeof(st));
      st.kind = g.kind;
      st.ia = col_index(g.a);
      st.ib = g.kind == 1 ? col_index(g.b) : 0;
      if (st.ia < 0 || st.ib < 0) return B200_OK;
      st.op1 = g.op1;
      st.op2 = g.op2;
      st.scalar = g.scalar;
      int found = -1;
      for (int q = 0; q < f.n_streams; ++q) {
        const FusedStream& o = f.streams[q];
        if (o.kind == st.kind && o.ia == st.ia && (st.kind == 0 || (o.ib == st.ib && o.op1 == st.op1 && o.op2 == st.op2 &&
                                                                    memcmp(&o.scalar, &st.scalar, sizeof(double)) == 0)))
          found = q;
      }
      if (found < 0) {
        if (f.n_streams == kFusedStreams) return B200_OK;
        f.streams[f.n_streams] = st;
        found = f.n_streams++;
      }
      a.s.aggs[j].col = found;
    }
    // (during pass 0 every new column comes with its own new plain stream, so stream i reads column i)
    if (pass == 0) f.n_plain = f.n_streams;
  }
  if (pred) {
    f.pred_data = reinterpret_cast<const char*>(pred->column.data) + pred->column.offset * 4;
    f.pred_type = pred->column.type;
    f.pred_op = pred->op;
    f.pred_scalar = pred->column.type == B200_INT32 ? int64_t(*reinterpret_cast<const int32_t*>(pred->scalar))
                                                    : int64_t(*reinterpret_cast<const uint32_t*>(pred->scalar));
  }
  const int grid = int(std::min<int64_t>(int64_t(ctx->num_sms) * 3, (n + 4095) / 4096));
  int64_t rpb = (n + grid - 1) / grid;
  rpb = (rpb + 31) / 32 * 32;
  a.rows_per_block = int(rpb);
  a.row_base = row_base;
  a.n_cols = f.n_streams;
  gb->row_bound = std::max(gb->row_bound, row_base + uint64_t(n));
  const uint32_t n_images = uint32_t(grid) * kTinyGroups;
  B200_RETURN_NOT_OK(gb_reserve_tmp(gb, size_t(n_images) * s.stride + 512));
  f.overflow = reinterpret_cast<uint32_t*>(gb->tmp);
  f.images = reinterpret_cast<char*>(gb->tmp) + 256;
  B200_CUDA(ctx, cudaMemsetAsync(f.overflow, 0, sizeof(uint32_t), ctx->stream));
  if (b200_ctx_once(ctx, 3)) {
    B200_CUDA(ctx, cudaFuncSetAttribute(gb_fused8_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(FusedSmem<1>))));
    B200_CUDA(ctx, cudaFuncSetAttribute(gb_fused8_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(FusedSmem<2>))));
    B200_CUDA(ctx, cudaFuncSetAttribute(gb_fused8_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(FusedSmem<3>))));
    B200_CUDA(ctx, cudaFuncSetAttribute(gb_fused8_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(sizeof(FusedSmem<4>))));
  }
  switch (f.n_streams) {
    case 1: B200_TIMED(ctx, "groupby_fused", (gb_fused8_kernel<1><<<grid, kTinyThreads, int(sizeof(FusedSmem<1>)), ctx->stream>>>(a, f))); break;
    case 2: B200_TIMED(ctx, "groupby_fused", (gb_fused8_kernel<2><<<grid, kTinyThreads, int(sizeof(FusedSmem<2>)), ctx->stream>>>(a, f))); break;
    case 3: B200_TIMED(ctx, "groupby_fused", (gb_fused8_kernel<3><<<grid, kTinyThreads, int(sizeof(FusedSmem<3>)), ctx->stream>>>(a, f))); break;
    default: B200_TIMED(ctx, "groupby_fused", (gb_fused8_kernel<4><<<grid, kTinyThreads, int(sizeof(FusedSmem<4>)), ctx->stream>>>(a, f))); break;
  }
  B200_CUDA(ctx, cudaMemcpyAsync(ctx->host_words + 24, f.overflow, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
  B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (*reinterpret_cast<const uint32_t*>(ctx->host_words + 24) != 0) return B200_OK;   // more than 8 groups: not taken
  B200_RETURN_NOT_OK(materialize(gb));
  gb_hp_import_kernel<<<(n_images + 255) / 256, 256, 0, ctx->stream>>>(gb->schema, f.images, n_images, gb->table, gb->counters);
  B200_LAUNCHED(ctx);
  uint32_t c[4];
  B200_RETURN_NOT_OK(read_counters(gb, c));
  *taken = true;
  return B200_OK;
}
}  // namespace

int b200_groupby_consume_fused(b200_groupby* gb, const b200_col* keys, const b200_fused_arg* args, uint64_t row_base,
                               const b200_predicate* pred) {
  B200_ENTER(gb ? gb->ctx : nullptr);
  if (!gb) return B200_INVALID;
  b200_ctx* ctx = gb->ctx;
  const GbSchema& s = gb->schema;
  if (!keys || (s.n_aggs > 0 && !args)) return b200_fail(ctx, B200_INVALID, "groupby_consume_fused: null argument");
  if (gb->dense_used) return b200_fail(ctx, B200_INVALID, "groupby: handle already used through b200_hash_agg_consume");
  gb->order_cache = nullptr;
  const int64_t n = keys[0].length;
  for (int j = 0; j < s.n_aggs; ++j) {
    const b200_fused_arg& g = args[j];
    if (g.kind != 0 && g.kind != 1) return b200_fail(ctx, B200_INVALID, "groupby_consume_fused: argument %d has kind %d", j, g.kind);
    if (g.kind == 1) {
      if (g.a.type != B200_DOUBLE || g.b.type != B200_DOUBLE)
        return b200_fail(ctx, B200_NOT_IMPLEMENTED, "groupby_consume_fused: derived arguments are float64 expressions");
      if (g.op1 < B200_ADD || g.op1 > B200_MULTIPLY || g.op2 < B200_ADD || g.op2 > B200_MULTIPLY)
        return b200_fail(ctx, B200_INVALID, "groupby_consume_fused: bad arithmetic operator");
      if (g.b.length != n) return b200_fail(ctx, B200_INVALID, "Array arguments must all be the same length");
    }
    if (g.a.length != n) return b200_fail(ctx, B200_INVALID, "Array arguments must all be the same length");
    const int want = g.kind == 1 ? B200_DOUBLE : g.a.type;
    if (want != s.aggs[j].arg_type) return b200_fail(ctx, B200_TYPE_ERROR, "groupby: argument %d has type %d, expected %d", j, want, s.aggs[j].arg_type);
  }
  if (pred) {
    if (!pred->scalar) return b200_fail(ctx, B200_INVALID, "groupby_consume_fused: predicate without a scalar");
    if (pred->column.length != n) return b200_fail(ctx, B200_INVALID, "Array arguments must all be the same length");
    if (pred->op < 0 || pred->op > 5) return b200_fail(ctx, B200_INVALID, "groupby_consume_fused: bad comparison %d", pred->op);
    if (col_has_nulls(pred->column)) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "groupby_consume_fused: predicate column with nulls (use Filter)");
  }
  if (n == 0) return B200_OK;
  bool taken = false;
  B200_RETURN_NOT_OK(try_fused8(gb, keys, args, row_base, pred, &taken));
  if (taken) return B200_OK;
  // ---- the same through the separate operations: comparison -> selection, arithmetic -> derived columns
  std::vector<void*> temps;
  auto release = [&]() {
    for (void* p : temps) b200_dev_free(ctx, p);
  };
  void* selection = nullptr;
  int st = B200_OK;
  if (pred) {
    st = b200_dev_alloc(ctx, &selection, size_t((n + 31) / 32) * 4);
    if (st == B200_OK) {
      temps.push_back(selection);
      st = b200_compare_scalar(ctx, &pred->column, pred->op, pred->scalar, selection);
    }
  }
  std::vector<b200_col> cols(std::max(s.n_aggs, 1));
  for (int j = 0; j < s.n_aggs && st == B200_OK; ++j) {
    const b200_fused_arg& g = args[j];
    if (g.kind == 0) {
      cols[j] = g.a;
      continue;
    }
    int same = -1;
    for (int q = 0; q < j; ++q)
      if (args[q].kind == 1 && same_col(args[q].a, g.a) && same_col(args[q].b, g.b) && args[q].op1 == g.op1 && args[q].op2 == g.op2 &&
          memcmp(&args[q].scalar, &g.scalar, sizeof(double)) == 0)
        same = q;
    if (same >= 0) {
      cols[j] = cols[same];
      continue;
    }
    void *inner = nullptr, *data = nullptr, *valid = nullptr;
    const bool nulls = col_has_nulls(g.a) || col_has_nulls(g.b);
    st = b200_dev_alloc(ctx, &inner, size_t(n) * 8);
    if (st == B200_OK) temps.push_back(inner);
    if (st == B200_OK) st = b200_dev_alloc(ctx, &data, size_t(n) * 8);
    if (st == B200_OK) temps.push_back(data);
    if (st == B200_OK && nulls) {
      st = b200_dev_alloc(ctx, &valid, size_t((n + 63) / 64) * 8);
      if (st == B200_OK) temps.push_back(valid);
    }
    if (st != B200_OK) break;
    st = b200_arithmetic(ctx, g.op2, nullptr, &g.scalar, &g.b, nullptr, inner, nullptr);   // validity of b rides with b below
    if (st != B200_OK) break;
    b200_col t = g.b;
    t.data = inner;
    t.offset = 0;
    if (col_has_nulls(g.b)) {
      // the inner result's validity is b's bitmap at b's offset: keep addressing it through b's descriptor
      t.validity = g.b.validity;
      t.offset = g.b.offset;
      t.data = reinterpret_cast<char*>(inner) - g.b.offset * 8;
    } else {
      t.validity = nullptr;
      t.null_count = 0;
    }
    st = b200_arithmetic(ctx, g.op1, &g.a, nullptr, &t, nullptr, data, valid);
    b200_col d;
    memset(&d, 0, sizeof(d));
    d.type = B200_DOUBLE;
    d.data = data;
    d.validity = valid;
    d.length = n;
    d.offset = 0;
    d.null_count = valid ? -1 : 0;
    cols[j] = d;
  }
  if (st == B200_OK) {
    st = selection ? b200_groupby_consume_selected(gb, keys, cols.data(), row_base, selection)
                   : b200_groupby_consume(gb, keys, cols.data(), row_base);
  }
  if (st == B200_OK) {
    // the temporaries go back to the context's block cache: make sure the kernels that read them have run
    cudaStreamSynchronize(ctx->stream);
  }
  release();
  return st;
}


// ---- Grouper::MakeGroupings (hash_aggregate.cc:968-1001): group ids -> per group the list of its rows ----------------
// The reference counts the ids, scans, and places row i at offsets[id[i]]++ (a stable counting sort).  Here: the ids
// are ranked by the stable radix sort (only the digits below num_groups are live), which IS the list values, and the
// offsets fall out of the sorted order — position i starts every group in (id of i-1, id of i].
namespace {
__global__ void mg_keys_kernel(const uint32_t* ids, int64_t n, uint32_t num_groups, uint64_t* keys, uint32_t* err) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t g = ids[i];
  if (g >= num_groups) atomicExch(err, 1u);
  keys[i] = g;
}
__global__ void mg_emit_kernel(const uint32_t* ids, const uint64_t* perm, int64_t n, uint32_t num_groups, int32_t* offsets,
                               int32_t* indices) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i > n) return;
  // position n closes the groups behind the last id
  const int64_t cur = i < n ? min(int64_t(ids[perm[i]]), int64_t(num_groups)) : int64_t(num_groups);   // (bad ids: reported, not written)
  const int64_t prev = i > 0 ? min(int64_t(ids[perm[i - 1]]), int64_t(num_groups)) : -1;
  for (int64_t g = prev + 1; g <= cur; ++g) offsets[g] = int32_t(i);
  if (i < n) indices[i] = int32_t(perm[i]);
}
}  // namespace

int b200_make_groupings_scratch_bytes(int64_t length, size_t* out_bytes) {
  if (!out_bytes || length < 0) return B200_INVALID;
  size_t sort_bytes = 0;
  b200_sort_pairs_scratch_bytes(length, &sort_bytes);
  *out_bytes = 2 * b200_align_up(size_t(length) * 8, 256) + sort_bytes + 1024;
  return B200_OK;
}

int b200_make_groupings(b200_ctx* ctx, const uint32_t* ids, int64_t length, uint32_t num_groups, int32_t* out_offsets,
                        int32_t* out_indices, void* scratch, size_t scratch_bytes) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (length < 0) return b200_fail(ctx, B200_INVALID, "make_groupings: negative length");
  if (!out_offsets) return b200_fail(ctx, B200_INVALID, "make_groupings: null offsets");
  // offsets and indices are int32 (list<int32>), like the reference's
  if (length > 0x7FFFFFFF) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "make_groupings: more than 2^31-1 rows");
  if (length == 0) {
    B200_CUDA(ctx, cudaMemsetAsync(out_offsets, 0, (size_t(num_groups) + 1) * 4, ctx->stream));
    return B200_OK;
  }
  size_t need = 0;
  b200_make_groupings_scratch_bytes(length, &need);
  if (!ids || !out_indices || !scratch || scratch_bytes < need) return b200_fail(ctx, B200_INVALID, "make_groupings: null argument or scratch too small");
  char* base = reinterpret_cast<char*>(b200_align_up(reinterpret_cast<uintptr_t>(scratch), 256));
  const size_t arr = b200_align_up(size_t(length) * 8, 256);
  uint64_t* keys = reinterpret_cast<uint64_t*>(base);
  uint64_t* perm = reinterpret_cast<uint64_t*>(base + arr);
  void* sort_scratch = base + 2 * arr;
  const size_t sort_bytes = scratch_bytes - size_t(base - reinterpret_cast<char*>(scratch)) - 2 * arr;
  const unsigned grid = unsigned((length + 255) / 256);
  mg_keys_kernel<<<grid, 256, 0, ctx->stream>>>(ids, length, num_groups, keys, ctx->async_err + 1);
  B200_LAUNCHED(ctx);
  int bits = 1;
  while (bits < 32 && (uint64_t(1) << bits) < uint64_t(num_groups)) ++bits;
  B200_RETURN_NOT_OK(b200_sort_pairs_bits(ctx, keys, nullptr, length, bits, perm, sort_scratch, sort_bytes));
  mg_emit_kernel<<<unsigned((length + 256) / 256), 256, 0, ctx->stream>>>(ids, perm, length, num_groups, out_offsets, out_indices);
  B200_LAUNCHED(ctx);
  // an id >= num_groups is a caller error (DCHECK upstream): reported, not written out of bounds
  B200_CUDA(ctx, cudaMemcpyAsync(ctx->host_words + 62, ctx->async_err + 1, sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
  B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (*reinterpret_cast<const uint32_t*>(ctx->host_words + 62) != 0) {
    B200_CUDA(ctx, cudaMemsetAsync(ctx->async_err + 1, 0, sizeof(uint32_t), ctx->stream));
    return b200_fail(ctx, B200_INDEX_ERROR, "make_groupings: group id >= num_groups (%u)", num_groups);
  }
  return B200_OK;
}

// ---- dictionary_encode's indices (DictEncodeAction, vector_hash.cc:165-226) from the Grouper's ids --------------------
// DictionaryEncodeOptions::MASK: the null key has no dictionary entry: its rows become null indices (value 0, as
// Int32Builder::UnsafeAppendNull leaves them) and every id handed out after the null group's shifts down by one.
namespace {
__global__ void dict_indices_kernel(const uint32_t* ids, int64_t n, uint32_t null_group, int32_t* out, uint32_t* validity) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  const bool in = i < n;
  const uint32_t id = in ? ids[i] : 0u;
  const bool valid = in && id != null_group;
  if (in) out[i] = valid ? int32_t(id - (id > null_group ? 1u : 0u)) : 0;
  const unsigned word = __ballot_sync(0xffffffffu, valid);
  if (validity && (threadIdx.x & 31) == 0 && in) validity[i >> 5] = word;
}
}  // namespace

int b200_dictionary_indices(b200_ctx* ctx, const uint32_t* ids, int64_t length, int64_t null_group, int32_t* out_indices,
                            void* out_validity) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (length < 0) return b200_fail(ctx, B200_INVALID, "dictionary_indices: negative length");
  if (length == 0) return B200_OK;
  if (!ids || !out_indices) return b200_fail(ctx, B200_INVALID, "dictionary_indices: null argument");
  if (null_group >= 0 && !out_validity) return b200_fail(ctx, B200_INVALID, "dictionary_indices: a null group needs a validity bitmap");
  if (null_group > int64_t(0xFFFFFFFEu)) return b200_fail(ctx, B200_INVALID, "dictionary_indices: null group out of range");
  dict_indices_kernel<<<unsigned((length + 255) / 256), 256, 0, ctx->stream>>>(
      ids, length, null_group < 0 ? 0xFFFFFFFFu : uint32_t(null_group), out_indices, reinterpret_cast<uint32_t*>(out_validity));
  B200_LAUNCHED(ctx);
  return B200_OK;
}

// ---- hash repartitioning -------------------------------------------------------------------------
int b200_partition_scratch_bytes(int64_t length, size_t* out_bytes) {
  if (!out_bytes || length < 0) return B200_INVALID;
  const size_t n_tiles = size_t((length + kPartTile - 1) / kPartTile);
  *out_bytes = b200_align_up(n_tiles * kPartMax * sizeof(uint32_t), 256) + 2 * kPartMax * sizeof(int64_t) + 512;
  return B200_OK;
}

int b200_groupby_partition_plan(b200_groupby* gb, const b200_col* keys, int n_parts, uint32_t* out_pos, int64_t* out_counts,
                                void* scratch, size_t scratch_bytes) {
  B200_ENTER(gb ? gb->ctx : nullptr);
  if (!gb) return B200_INVALID;
  b200_ctx* ctx = gb->ctx;
  if (!keys || !out_counts) return b200_fail(ctx, B200_INVALID, "partition_plan: null argument");
  if (n_parts < 1 || n_parts > kPartMax) return b200_fail(ctx, B200_INVALID, "partition_plan: 1..%d partitions", kPartMax);
  if (gb->wide) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "partition_plan: keys wider than 64 packed bits");
  GbArgs a;
  memset(&a, 0, sizeof(a));
  a.s = gb->schema;
  B200_RETURN_NOT_OK(fill_keys(gb, keys, &a));
  for (int p = 0; p < n_parts; ++p) out_counts[p] = 0;
  if (a.n == 0) return B200_OK;
  if (a.n >= (int64_t(1) << 32)) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "partition_plan: more than 2^32-1 rows per call");
  size_t need = 0;
  b200_partition_scratch_bytes(a.n, &need);
  if (!out_pos || !scratch || scratch_bytes < need) return b200_fail(ctx, B200_INVALID, "partition_plan: scratch too small (%zu < %zu)", scratch_bytes, need);
  const int64_t n_tiles = (a.n + kPartTile - 1) / kPartTile;
  char* base = reinterpret_cast<char*>(b200_align_up(reinterpret_cast<uintptr_t>(scratch), 256));
  uint32_t* tile_counts = reinterpret_cast<uint32_t*>(base);
  int64_t* totals = reinterpret_cast<int64_t*>(base + b200_align_up(size_t(n_tiles) * kPartMax * sizeof(uint32_t), 256));
  B200_TIMED(ctx, "partition_plan", (gb_part_count_kernel<<<(unsigned)n_tiles, kPartThreads, 0, ctx->stream>>>(a, uint32_t(n_parts), tile_counts)));
  gb_part_scan_kernel<<<1, 1024, 0, ctx->stream>>>(tile_counts, n_tiles, uint32_t(n_parts), totals);
  B200_LAUNCHED(ctx);
  B200_TIMED(ctx, "partition_plan", (gb_part_pos_kernel<<<(unsigned)n_tiles, kPartThreads, 0, ctx->stream>>>(a, uint32_t(n_parts), tile_counts, totals, out_pos)));
  B200_CUDA(ctx, cudaMemcpyAsync(ctx->host_words + 32, totals, kPartMax * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  for (int p = 0; p < n_parts; ++p) out_counts[p] = ctx->host_words[32 + p];
  return B200_OK;
}

int b200_partition_scatter(b200_ctx* ctx, const b200_col* col, const uint32_t* pos, void* out_data, uint8_t* out_valid) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (!col) return b200_fail(ctx, B200_INVALID, "partition_scatter: null column");
  const int64_t n = col->length;
  if (n == 0) return B200_OK;
  if (!pos || !out_data || !col->data) return b200_fail(ctx, B200_INVALID, "partition_scatter: null argument");
  const int bits = b200_type_bits(col->type);
  if (bits < 8) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "partition_scatter: type %d", col->type);
  const bool has_nulls = col->validity != nullptr && col->null_count != 0;
  const BitmapRef vv = has_nulls ? make_bitmap_ref(col->validity, col->offset) : BitmapRef{nullptr, 0};
  const char* data = reinterpret_cast<const char*>(col->data) + col->offset * (bits / 8);
  const unsigned grid = unsigned((n + 255) / 256);
  switch (bits) {
    case 8: B200_TIMED(ctx, "partition_scatter", (gb_part_scatter_kernel<uint8_t><<<grid, 256, 0, ctx->stream>>>(reinterpret_cast<const uint8_t*>(data), vv, pos, n, reinterpret_cast<uint8_t*>(out_data), out_valid))); break;
    case 16: B200_TIMED(ctx, "partition_scatter", (gb_part_scatter_kernel<uint16_t><<<grid, 256, 0, ctx->stream>>>(reinterpret_cast<const uint16_t*>(data), vv, pos, n, reinterpret_cast<uint16_t*>(out_data), out_valid))); break;
    case 32: B200_TIMED(ctx, "partition_scatter", (gb_part_scatter_kernel<uint32_t><<<grid, 256, 0, ctx->stream>>>(reinterpret_cast<const uint32_t*>(data), vv, pos, n, reinterpret_cast<uint32_t*>(out_data), out_valid))); break;
    default: B200_TIMED(ctx, "partition_scatter", (gb_part_scatter_kernel<uint64_t><<<grid, 256, 0, ctx->stream>>>(reinterpret_cast<const uint64_t*>(data), vv, pos, n, reinterpret_cast<uint64_t*>(out_data), out_valid))); break;
  }
  return B200_OK;
}

int b200_partition_rows(b200_ctx* ctx, int64_t length, uint64_t row_base, const uint32_t* pos, uint64_t* out_rows) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (length < 0) return b200_fail(ctx, B200_INVALID, "partition_rows: negative length");
  if (length == 0) return B200_OK;
  if (!pos || !out_rows) return b200_fail(ctx, B200_INVALID, "partition_rows: null argument");
  gb_part_rows_kernel<<<unsigned((length + 255) / 256), 256, 0, ctx->stream>>>(length, row_base, pos, out_rows);
  B200_LAUNCHED(ctx);
  return B200_OK;
}

int b200_pack_validity(b200_ctx* ctx, const uint8_t* valid_bytes, int64_t length, void* out_bitmap, int64_t* out_null_count) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (length < 0) return b200_fail(ctx, B200_INVALID, "pack_validity: negative length");
  if (out_null_count) *out_null_count = 0;
  if (length == 0) return B200_OK;
  if (!valid_bytes || !out_bitmap) return b200_fail(ctx, B200_INVALID, "pack_validity: null argument");
  B200_RETURN_NOT_OK(b200_ctx_reserve_scratch(ctx, 256));
  unsigned long long* counter = reinterpret_cast<unsigned long long*>(ctx->scratch);
  B200_CUDA(ctx, cudaMemsetAsync(counter, 0, sizeof(unsigned long long), ctx->stream));
  gb_pack_validity_kernel<<<unsigned((length + 255) / 256), 256, 0, ctx->stream>>>(valid_bytes, length, reinterpret_cast<uint32_t*>(out_bitmap), counter);
  B200_LAUNCHED(ctx);
  if (out_null_count) {
    B200_CUDA(ctx, cudaMemcpyAsync(ctx->host_words + 48, counter, sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
    B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *out_null_count = ctx->host_words[48];
  }
  return B200_OK;
}

int b200_groupby_num_groups(b200_groupby* gb, uint32_t* out) {
  B200_ENTER(gb ? gb->ctx : nullptr);
  if (!gb || !out) return B200_INVALID;
  if (gb->dense_used) {
    *out = gb->dense_groups;
    return B200_OK;
  }
  uint32_t c[4];
  B200_RETURN_NOT_OK(read_counters(gb, c));
  *out = c[0];
  return B200_OK;
}

int b200_grouper_consume(b200_groupby* gb, const b200_col* keys, uint32_t* out_ids) {
  B200_ENTER(gb ? gb->ctx : nullptr);
  if (!gb) return B200_INVALID;
  b200_ctx* ctx = gb->ctx;
  if (!keys) return b200_fail(ctx, B200_INVALID, "grouper_consume: null keys");
  gb->order_cache = nullptr;
  B200_NARROW_KEYS(gb, keys, nullptr);
  GbArgs a;
  memset(&a, 0, sizeof(a));
  a.s = gb->schema;
  a.s.n_aggs = 0;
  a.s.lstride = 16;
  B200_RETURN_NOT_OK(fill_keys(gb, keys, &a));
  const int64_t n = a.n;
  if (n == 0) return B200_OK;
  if (!out_ids) return b200_fail(ctx, B200_INVALID, "grouper_consume: null output");
  B200_RETURN_NOT_OK(materialize(gb));
  // row indices continue across calls so that first-occurrence order is global
  a.row_base = gb->dense_rows;
  gb->dense_rows += uint64_t(n);
  gb->row_bound = std::max(gb->row_bound, gb->dense_rows);
  // make room up front: a rehash in the middle of a batch would invalidate positional slot ids
  {
    uint32_t c[4];
    B200_RETURN_NOT_OK(read_counters(gb, c));
    if ((uint64_t(gb->groups) + uint64_t(n)) * 2 > gb->table.cap) B200_RETURN_NOT_OK(grow_table(gb, (uint64_t(gb->groups) + uint64_t(n)) * 2));
  }
  // all scratch of this call is reserved up front: the claimed-slot list written by pass 1 must
  // survive until the ids are assigned
  size_t sort_bytes = 0;
  b200_sort_pairs_scratch_bytes(n, &sort_bytes);
  const size_t words = size_t((n + 31) / 32);
  const size_t base_off = 2 * b200_align_up(words * 4, 256) + b200_align_up(size_t(n) * 4, 256) + 256;
  const size_t arr = b200_align_up(size_t(n) * 8, 256);
  B200_RETURN_NOT_OK(gb_reserve_tmp(gb, base_off + 3 * arr + sort_bytes + 256));
  const uint64_t n_slots = gb->table.cap + 2;
  const bool rebuild_ids = gb->slot_ids == nullptr;
  if (rebuild_ids) {
    B200_RETURN_NOT_OK(b200_dev_alloc(ctx, reinterpret_cast<void**>(&gb->slot_ids), n_slots * sizeof(uint32_t)));
    gb->slot_ids_cap = n_slots;
    gb_fill_u32_kernel<<<(unsigned)((n_slots + 255) / 256), 256, 0, ctx->stream>>>(gb->slot_ids, n_slots, kNoId);
    B200_LAUNCHED(ctx);
    gb->ids_assigned = 0;
  }
  // pass 1: insert keys, track first rows and newly claimed slots
  a.mode = kModeInsert;
  a.slot_ids = gb->slot_ids;
  B200_RETURN_NOT_OK(run_consume(gb, a));
  uint32_t c[4];
  B200_RETURN_NOT_OK(read_counters(gb, c));
  // slots that still need an id: the claimed list, or (after a rehash / fused consumes) a full scan
  uint32_t n_new = c[2];
  const bool full_scan = rebuild_ids && gb->groups != n_new;
  const uint32_t want = gb->groups - gb->ids_assigned;
  if (want > 0) {
    if (want > uint64_t(n)) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "grouper: %u groups from earlier fused consumes need ids; use a fresh handle", want);
    // scratch: first_rows[n] slot_idx[n] sorted[n] + sort scratch, placed after run_consume's regions
    char* base = reinterpret_cast<char*>(gb->tmp);
    uint32_t* claimed = reinterpret_cast<uint32_t*>(base + 2 * b200_align_up(words * 4, 256));
    uint64_t* first_rows = reinterpret_cast<uint64_t*>(base + base_off);
    uint64_t* slot_idx = reinterpret_cast<uint64_t*>(base + base_off + arr);
    uint64_t* sorted = reinterpret_cast<uint64_t*>(base + base_off + 2 * arr);
    void* sort_scratch = base + base_off + 3 * arr;
    if (full_scan || n_new != want) {
      B200_CUDA(ctx, cudaMemsetAsync(gb->counters + 4, 0, sizeof(uint32_t), ctx->stream));
      gb_collect_kernel<<<(unsigned)((n_slots + 255) / 256), 256, 0, ctx->stream>>>(gb->table, gb->schema.stride, first_rows, slot_idx,
                                                                                   gb->counters + 4, gb->slot_ids, 1);
      B200_LAUNCHED(ctx);
    } else {
      gb_claimed_rows_kernel<<<(want + 255) / 256, 256, 0, ctx->stream>>>(gb->table, gb->schema.stride, claimed, want, first_rows, slot_idx);
      B200_LAUNCHED(ctx);
    }
    B200_RETURN_NOT_OK(b200_sort_pairs(ctx, first_rows, slot_idx, want, nullptr, sorted, sort_scratch, sort_bytes));
    gb_assign_ids_kernel<<<(want + 255) / 256, 256, 0, ctx->stream>>>(sorted, want, gb->ids_assigned, gb->slot_ids);
    B200_LAUNCHED(ctx);
    gb->ids_assigned += want;
  }
  // pass 2: look the ids up
  a.mode = kModeLookup;
  a.out_ids = out_ids;
  a.claimed = nullptr;
  a.row_base = 0;
  B200_RETURN_NOT_OK(run_consume(gb, a));
  return B200_OK;
}

int b200_hash_agg_consume(b200_groupby* gb, int agg_index, const b200_col* arg, const uint32_t* ids, uint32_t num_groups) {
  B200_ENTER(gb ? gb->ctx : nullptr);
  if (!gb) return B200_INVALID;
  b200_ctx* ctx = gb->ctx;
  if (agg_index < 0 || agg_index >= gb->schema.n_aggs) return b200_fail(ctx, B200_INVALID, "hash_agg_consume: bad aggregate index");
  if (!arg || (arg->length > 0 && !ids)) return b200_fail(ctx, B200_INVALID, "hash_agg_consume: null argument");
  if (arg->type != gb->schema.aggs[agg_index].arg_type) return b200_fail(ctx, B200_TYPE_ERROR, "hash_agg_consume: argument type mismatch");
  gb->dense_used = true;
  // MaybeReserve (hash_aggregate.cc:450-458): grow the per-group state to num_groups
  if (num_groups > gb->dense_cap) {
    const uint32_t new_cap = uint32_t(pow2_at_least(num_groups));
    char* nd = nullptr;
    B200_RETURN_NOT_OK(b200_dev_alloc(ctx, reinterpret_cast<void**>(&nd), size_t(new_cap) * gb->schema.stride));
    gb_init_kernel<<<(new_cap + 255) / 256, 256, 0, ctx->stream>>>(gb->schema, nd, new_cap, gb->schema.stride);
    B200_LAUNCHED(ctx);
    if (gb->dense) {
      B200_CUDA(ctx, cudaMemcpyAsync(nd, gb->dense, size_t(gb->dense_cap) * gb->schema.stride, cudaMemcpyDeviceToDevice, ctx->stream));
      B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
      b200_dev_free(ctx, gb->dense);
    }
    gb->dense = nd;
    gb->dense_cap = new_cap;
  }
  if (num_groups > gb->dense_groups) gb->dense_groups = num_groups;
  const int64_t n = arg->length;
  if (n == 0) return B200_OK;
  GbCol col;
  fill_col(*arg, &col);
  gb_dense_consume_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(gb->schema, agg_index, col, ids, n, gb->dense,
                                                                               num_groups, gb->counters);
  B200_LAUNCHED(ctx);
  return B200_OK;
}

int b200_hash_agg_merge(b200_groupby* dst, b200_groupby* src) {
  B200_ENTER(dst ? dst->ctx : nullptr);
  if (!dst || !src) return B200_INVALID;
  b200_ctx* ctx = dst->ctx;
  if (dst->ctx != src->ctx) return b200_fail(ctx, B200_INVALID, "hash_agg_merge: handles belong to different contexts");
  if (memcmp(&dst->schema, &src->schema, sizeof(GbSchema)) != 0) return b200_fail(ctx, B200_TYPE_ERROR, "hash_agg_merge: schemas differ");
  if (!src->dense_used || src->dense_groups == 0) return B200_OK;
  dst->dense_used = true;
  const uint32_t n = src->dense_groups;
  if (n > dst->dense_cap) {  // MaybeReserve, as in b200_hash_agg_consume
    const uint32_t new_cap = uint32_t(pow2_at_least(n));
    char* nd = nullptr;
    B200_RETURN_NOT_OK(b200_dev_alloc(ctx, reinterpret_cast<void**>(&nd), size_t(new_cap) * dst->schema.stride));
    gb_init_kernel<<<(new_cap + 255) / 256, 256, 0, ctx->stream>>>(dst->schema, nd, new_cap, dst->schema.stride);
    B200_LAUNCHED(ctx);
    if (dst->dense) {
      B200_CUDA(ctx, cudaMemcpyAsync(nd, dst->dense, size_t(dst->dense_cap) * dst->schema.stride, cudaMemcpyDeviceToDevice, ctx->stream));
      B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
      b200_dev_free(ctx, dst->dense);
    }
    dst->dense = nd;
    dst->dense_cap = new_cap;
  }
  if (n > dst->dense_groups) dst->dense_groups = n;
  gb_dense_merge_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(dst->schema, src->dense, dst->dense, n);
  B200_LAUNCHED(ctx);
  return B200_OK;
}

int b200_groupby_state_layout(b200_groupby* gb, size_t* slot_bytes) {
  B200_ENTER(gb ? gb->ctx : nullptr);
  if (!gb || !slot_bytes) return B200_INVALID;
  *slot_bytes = size_t(gb->schema.stride);
  return B200_OK;
}

static int collect_slots(b200_groupby* gb, uint64_t** first_rows, uint64_t** slot_idx, uint32_t* n_out, size_t extra,
                         char** extra_ptr) {
  b200_ctx* ctx = gb->ctx;
  uint32_t c[4];
  B200_RETURN_NOT_OK(read_counters(gb, c));
  const uint32_t G = c[0];
  const size_t arr = b200_align_up(size_t(std::max<uint32_t>(G, 1)) * 8, 256);
  B200_RETURN_NOT_OK(gb_reserve_tmp(gb, 2 * arr + extra + 256));
  *first_rows = reinterpret_cast<uint64_t*>(gb->tmp);
  *slot_idx = reinterpret_cast<uint64_t*>(reinterpret_cast<char*>(gb->tmp) + arr);
  if (extra_ptr) *extra_ptr = reinterpret_cast<char*>(gb->tmp) + 2 * arr;
  *n_out = G;
  if (gb->pending) {
    if (G > 0) {
      gb_collect_images_kernel<<<(G + 255) / 256, 256, 0, ctx->stream>>>(gb->pending, gb->schema.stride, G, *first_rows, *slot_idx);
      B200_LAUNCHED(ctx);
    }
    return B200_OK;
  }
  B200_RETURN_NOT_OK(ensure_table(gb));
  B200_CUDA(ctx, cudaMemsetAsync(gb->counters + 4, 0, sizeof(uint32_t), ctx->stream));
  const uint64_t n_slots = gb->table.cap + 2;
  gb_collect_kernel<<<(unsigned)((n_slots + 255) / 256), 256, 0, ctx->stream>>>(gb->table, gb->schema.stride, *first_rows, *slot_idx,
                                                                               gb->counters + 4, nullptr, 0);
  B200_LAUNCHED(ctx);
  return B200_OK;
}

int b200_groupby_first_rows(b200_groupby* gb, uint64_t* out_first_rows) {
  B200_ENTER(gb ? gb->ctx : nullptr);
  if (!gb) return B200_INVALID;
  b200_ctx* ctx = gb->ctx;
  if (gb->dense_used) return b200_fail(ctx, B200_INVALID, "groupby_first_rows: dense handle has no row order");
  if (gb->order_cache && out_first_rows) {   // finalize has just ranked the groups
    if (gb->order_n) {
      gb_first_rows_kernel<<<(gb->order_n + 255) / 256, 256, 0, ctx->stream>>>(gb->order_slots, gb->schema.stride, gb->order_cache,
                                                                                gb->order_n, out_first_rows);
      B200_LAUNCHED(ctx);
    }
    return B200_OK;
  }
  uint64_t *first_rows, *slot_idx;
  size_t sort_bytes = 0;
  uint32_t c[4];
  B200_RETURN_NOT_OK(read_counters(gb, c));
  b200_sort_pairs_scratch_bytes(std::max<uint32_t>(c[0], 1), &sort_bytes);
  const size_t arr = b200_align_up(size_t(std::max<uint32_t>(c[0], 1)) * 8, 256);
  char* extra = nullptr;
  uint32_t G = 0;
  B200_RETURN_NOT_OK(collect_slots(gb, &first_rows, &slot_idx, &G, arr + sort_bytes + 256, &extra));
  if (G == 0) return B200_OK;
  if (!out_first_rows) return b200_fail(ctx, B200_INVALID, "groupby_first_rows: null output");
  return b200_sort_pairs(ctx, first_rows, slot_idx, G, out_first_rows, reinterpret_cast<uint64_t*>(extra), extra + arr, sort_bytes);
}

int b200_groupby_export(b200_groupby* gb, void* out_state, uint32_t* out_groups) {
  B200_ENTER(gb ? gb->ctx : nullptr);
  if (!gb || !out_groups) return B200_INVALID;
  b200_ctx* ctx = gb->ctx;
  uint64_t *first_rows, *slot_idx;
  uint32_t G;
  if (gb->wide) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "groupby_export: keys wider than 64 packed bits (the key store is not part of the state image)");
  B200_RETURN_NOT_OK(materialize(gb));
  B200_RETURN_NOT_OK(collect_slots(gb, &first_rows, &slot_idx, &G, 0, nullptr));
  if (G > 0) {
    if (!out_state) return b200_fail(ctx, B200_INVALID, "groupby_export: null output");
    gb_export_kernel<<<(G + 255) / 256, 256, 0, ctx->stream>>>(gb->table, gb->schema.stride, slot_idx, G,
                                                               reinterpret_cast<char*>(out_state), nullptr);
    B200_LAUNCHED(ctx);
  }
  B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *out_groups = G;
  return B200_OK;
}

int b200_groupby_import(b200_groupby* gb, const void* state, uint32_t n_slots) {
  B200_ENTER(gb ? gb->ctx : nullptr);
  if (!gb) return B200_INVALID;
  b200_ctx* ctx = gb->ctx;
  if (n_slots == 0) return B200_OK;
  if (!state) return b200_fail(ctx, B200_INVALID, "groupby_import: null state");
  gb->order_cache = nullptr;
  if (gb->wide) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "groupby_import: keys wider than 64 packed bits");
  gb->row_bound = ~uint64_t(0);
  B200_RETURN_NOT_OK(materialize(gb));
  uint32_t c[4];
  B200_RETURN_NOT_OK(read_counters(gb, c));
  B200_RETURN_NOT_OK(grow_table(gb, 2 * (uint64_t(gb->groups) + n_slots)));
  gb_import_kernel<<<(n_slots + 255) / 256, 256, 0, ctx->stream>>>(gb->schema, reinterpret_cast<const char*>(state), n_slots,
                                                                  gb->table, gb->counters);
  B200_LAUNCHED(ctx);
  return read_counters(gb, c);
}

int b200_groupby_merge(b200_groupby* dst, b200_groupby* src) {
  B200_ENTER(dst ? dst->ctx : nullptr);
  if (!dst || !src) return B200_INVALID;
  b200_ctx* ctx = dst->ctx;
  if (memcmp(&dst->schema, &src->schema, sizeof(GbSchema)) != 0) return b200_fail(ctx, B200_TYPE_ERROR, "groupby_merge: schemas differ");
  if (dst->ctx != src->ctx) return b200_fail(ctx, B200_INVALID, "groupby_merge: handles belong to different contexts");
  dst->order_cache = nullptr;
  if (dst->wide || src->wide) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "groupby_merge: keys wider than 64 packed bits");
  dst->row_bound = std::max(dst->row_bound, src->row_bound);
  B200_RETURN_NOT_OK(materialize(dst));
  B200_RETURN_NOT_OK(materialize(src));
  uint32_t c[4];
  B200_RETURN_NOT_OK(read_counters(src, c));
  const uint32_t src_groups = c[0];
  B200_RETURN_NOT_OK(read_counters(dst, c));
  B200_RETURN_NOT_OK(grow_table(dst, 2 * (uint64_t(dst->groups) + src_groups)));
  const uint64_t n_src = src->table.cap + 2;
  gb_move_kernel<<<(unsigned)((n_src + 255) / 256), 256, 0, ctx->stream>>>(dst->schema, src->table.slots, n_src, src->schema.stride,
                                                                           src->table.cap, dst->table, 1, dst->counters);
  B200_LAUNCHED(ctx);
  return read_counters(dst, c);
}

int b200_groupby_finalize(b200_groupby* gb, b200_out_col* out_aggs, b200_out_col* out_keys) {
  B200_ENTER(gb ? gb->ctx : nullptr);
  if (!gb) return B200_INVALID;
  b200_ctx* ctx = gb->ctx;
  EmitArgs e;
  memset(&e, 0, sizeof(e));
  e.s = gb->schema;
  uint32_t G = 0;
  if (gb->dense_used) {
    // HashAggregateKernel::finalize only: ids are already dense, no keys to emit
    uint32_t c[4];
    B200_RETURN_NOT_OK(read_counters(gb, c));
    G = gb->dense_groups;
    e.t.slots = gb->dense;
    e.t.cap = 0;
    e.order = nullptr;
    B200_RETURN_NOT_OK(gb_reserve_tmp(gb, 4096));
    e.null_counts = reinterpret_cast<uint32_t*>(gb->tmp);
  } else {
    uint64_t *first_rows, *slot_idx;
    size_t sort_bytes = 0;
    uint32_t c[4];
    B200_RETURN_NOT_OK(read_counters(gb, c));
    b200_sort_pairs_scratch_bytes(std::max<uint32_t>(c[0], 1), &sort_bytes);
    const size_t arr = b200_align_up(size_t(std::max<uint32_t>(c[0], 1)) * 8, 256);
    char* extra = nullptr;
    B200_RETURN_NOT_OK(collect_slots(gb, &first_rows, &slot_idx, &G, arr + sort_bytes + 4096, &extra));
    uint64_t* sorted = reinterpret_cast<uint64_t*>(extra);
    e.null_counts = reinterpret_cast<uint32_t*>(extra + arr);
    void* sort_scratch = extra + arr + 4096;
    // pending images: slot index == position, so the sort emits the permutation itself (no gather of a payload)
    const int row_bits = gb->row_bound <= (uint64_t(1) << 32) ? 32 : 64;
    if (G > 0)
      B200_RETURN_NOT_OK(b200_sort_pairs_bits(ctx, first_rows, gb->pending ? nullptr : slot_idx, G, row_bits, sorted, sort_scratch, sort_bytes));
    e.t = gb->table;
    if (gb->pending) {   // slot index = image index; no image is a special slot
      e.t.slots = gb->pending;
      e.t.cap = ~uint64_t(0) - 16;   // cap and cap + 1 (the special slots) match no image index
    }
    e.order = sorted;
    gb->order_cache = sorted;
    gb->order_slots = e.t.slots;
    gb->order_n = G;
  }
  e.n_groups = G;
  B200_CUDA(ctx, cudaMemsetAsync(e.null_counts, 0, (kMaxKeys + 2 * kMaxAggs) * sizeof(uint32_t), ctx->stream));
  uint32_t* wide_ids = nullptr;   // wide keys: the stand-in uint32 key column is emitted here, then decoded
  if (gb->wide && out_keys && !gb->dense_used) {
    B200_RETURN_NOT_OK(b200_dev_alloc(ctx, reinterpret_cast<void**>(&wide_ids), size_t(std::max<uint32_t>(G, 1)) * 4 + 512));
    e.keys[0].data = wide_ids;
    e.keys[0].validity = nullptr;
    e.keys[0].type = B200_UINT32;
    for (int k = 0; k < gb->wn_keys; ++k)
      if (G > 0 && !out_keys[k].data) {
        b200_dev_free(ctx, wide_ids);
        return b200_fail(ctx, B200_INVALID, "groupby_finalize: key %d has no data buffer", k);
      }
  }
  for (int k = 0; k < e.s.n_keys && !gb->wide; ++k) {
    if (out_keys && !gb->dense_used) {
      e.keys[k].data = out_keys[k].data;
      e.keys[k].validity = reinterpret_cast<uint32_t*>(out_keys[k].validity);
      e.keys[k].type = e.s.key_type[k];
      if (G > 0 && !out_keys[k].data) return b200_fail(ctx, B200_INVALID, "groupby_finalize: key %d has no data buffer", k);
    }
  }
  int oi = 0;
  for (int j = 0; j < e.s.n_aggs; ++j) {
    const int n_out = e.s.aggs[j].kind == B200_HASH_MIN_MAX ? 2 : 1;
    for (int q = 0; q < n_out; ++q, ++oi) {
      if (!out_aggs) return b200_fail(ctx, B200_INVALID, "groupby_finalize: null out_aggs");
      e.aggs[oi].data = out_aggs[oi].data;
      e.aggs[oi].validity = reinterpret_cast<uint32_t*>(out_aggs[oi].validity);
      if (G > 0 && !out_aggs[oi].data) return b200_fail(ctx, B200_INVALID, "groupby_finalize: aggregate output %d has no data buffer", oi);
    }
  }
  if (G > 0) {
    gb_emit_kernel<<<(G + 255) / 256, 256, 0, ctx->stream>>>(e);
    B200_LAUNCHED(ctx);
  }
  if (wide_ids) {
    WideDecodeArgs d;
    memset(&d, 0, sizeof(d));
    d.ids = wide_ids;
    d.n_groups = G;
    d.store = gb->wstore;
    d.words = gb->wwords;
    d.n_keys = gb->wn_keys;
    d.null_counts = wide_ids + ((size_t(std::max<uint32_t>(G, 1)) * 4 + 255) / 256 * 256) / 4;
    d.n_var = gb->wn_var;
    d.store_var = gb->wstore_var;
    d.heap = gb->wheap;
    for (int k = 0; k < gb->wn_keys; ++k) {
      d.nbits[k] = gb->wbits[k];
      d.byte_off[k] = gb->woff[k];
      d.data[k] = out_keys[k].data;
      d.validity[k] = reinterpret_cast<uint32_t*>(out_keys[k].validity);
      d.var_index[k] = gb->wvar_index[k];
      d.out_bytes[k] = reinterpret_cast<uint8_t*>(out_keys[k].extra);
      if (gb->wtypes[k] == B200_BINARY && G > 0 && gb->wvar_bytes_host[gb->wvar_index[k]] > 0 && !out_keys[k].extra) {
        b200_dev_free(ctx, wide_ids);
        return b200_fail(ctx, B200_INVALID, "groupby_finalize: variable-length key %d has no value buffer (b200_groupby_key_bytes)", k);
      }
    }
    B200_CUDA(ctx, cudaMemsetAsync(d.null_counts, 0, kWideMaxKeys * sizeof(uint32_t), ctx->stream));
    if (G > 0) {
      wide_decode_kernel<<<(G + 255) / 256, 256, 0, ctx->stream>>>(d);
      B200_LAUNCHED(ctx);
    }
    // variable-length keys: the lengths just written become the int32 offsets (one scan per column), then the bytes
    for (int k = 0; k < gb->wn_keys; ++k) {
      if (gb->wtypes[k] != B200_BINARY) continue;
      uint32_t* off = reinterpret_cast<uint32_t*>(out_keys[k].data);
      if (G == 0) {
        if (off) B200_CUDA(ctx, cudaMemsetAsync(off, 0, sizeof(uint32_t), ctx->stream));
        continue;
      }
      gb_hp_scan_kernel<<<1, 1024, 0, ctx->stream>>>(off, G, nullptr, off);
      B200_LAUNCHED(ctx);
    }
    if (G > 0 && gb->wn_var) {
      wide_decode_bytes_kernel<<<(G + 255) / 256, 256, 0, ctx->stream>>>(d);
      B200_LAUNCHED(ctx);
    }
    B200_CUDA(ctx, cudaMemcpyAsync(ctx->host_words + 44, d.null_counts, kWideMaxKeys * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
  }
  B200_CUDA(ctx, cudaMemcpyAsync(ctx->host_words + 24, e.null_counts, (kMaxKeys + 2 * kMaxAggs) * sizeof(uint32_t),
                                 cudaMemcpyDeviceToHost, ctx->stream));
  B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  const uint32_t* nc = reinterpret_cast<const uint32_t*>(ctx->host_words + 24);
  if (wide_ids) {
    const uint32_t* wnc = reinterpret_cast<const uint32_t*>(ctx->host_words + 44);
    for (int k = 0; k < gb->wn_keys; ++k) {
      out_keys[k].length = G;
      out_keys[k].null_count = wnc[k];
      out_keys[k].type = gb->wtypes[k];
    }
    b200_dev_free(ctx, wide_ids);
  }
  for (int k = 0; k < e.s.n_keys && out_keys && !gb->dense_used && !gb->wide; ++k) {
    out_keys[k].length = G;
    out_keys[k].null_count = nc[k];
    out_keys[k].type = e.s.key_type[k];
  }
  for (int i = 0; i < oi; ++i) {
    out_aggs[i].length = G;
    out_aggs[i].null_count = nc[kMaxKeys + i];
  }
  return B200_OK;
}

}  // extern "C"
