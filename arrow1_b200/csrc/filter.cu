// Filter: order-preserving stream compaction of a fixed-width column by a boolean mask, with
// validity propagation.  Replaces GetFilterOutputSize + PrimitiveFilter + PrimitiveFilterImpl
// (reference compute/kernels/vector_selection.cc:61-88, 570-835).
//
// Two kernels:
//   filter_plan_kernel   popcount scan of the mask (∧ / ∨¬ its validity): per-tile exclusive output
//                        offsets (4096-row tiles grouped in 64-tile super-tiles) and the output
//                        length; the last block to finish scans the super-tile totals.
//   filter_apply_kernel  one 4096-row tile per CTA; warp ballots are replaced by the mask words
//                        themselves (the mask IS the ballot), ranks are popc(mask & lanemask_lt),
//                        values are gathered with predicated coalesced loads and written to
//                        consecutive output slots; validity bits (and boolean values) are
//                        compacted with REDUX.OR and appended as whole 32-bit words.
// HBM roofline: 8 B/row values + 3/8 B/row bitmaps read, s*(8+1/8) B/row written.
#include <type_traits>

#include <vector>
#include <mutex>

#include "common.cuh"

namespace {

constexpr int kTileRows = 4096;       // rows per apply CTA
constexpr int kTilesPerSuper = 64;    // tiles per plan CTA
constexpr int kSuperRows = kTileRows * kTilesPerSuper;
constexpr int kThreads = 256;

struct PlanHeader {       // first 64 bytes of a plan
  int64_t out_length;     // total number of emitted rows
  uint32_t done_counter;  // blocks finished (threadfence reduction)
  uint32_t pad;
  int64_t reserved[6];
};
static_assert(sizeof(PlanHeader) == 64, "plan header");

struct PlanView {
  PlanHeader* header;
  int64_t* super_prefix;   // [n_super] exclusive prefix of emitted rows per super-tile
  int64_t* super_total;    // [n_super]
  uint32_t* tile_prefix;   // [n_tiles] exclusive prefix inside the super-tile
};

inline int64_t n_tiles_for(int64_t length) { return (length + kTileRows - 1) / kTileRows; }
inline int64_t n_super_for(int64_t length) { return (length + kSuperRows - 1) / kSuperRows; }

inline size_t plan_bytes_for(int64_t length) {
  int64_t nt = n_tiles_for(length), ns = n_super_for(length);
  return sizeof(PlanHeader) + b200_align_up(size_t(ns) * 16, 64) + b200_align_up(size_t(nt) * 4, 64) + 64;
}

inline PlanView plan_view(void* plan, int64_t length) {
  PlanView v;
  int64_t ns = n_super_for(length);
  char* p = reinterpret_cast<char*>(plan);
  v.header = reinterpret_cast<PlanHeader*>(p);
  v.super_prefix = reinterpret_cast<int64_t*>(p + sizeof(PlanHeader));
  v.super_total = v.super_prefix + ns;
  v.tile_prefix = reinterpret_cast<uint32_t*>(p + sizeof(PlanHeader) + b200_align_up(size_t(ns) * 16, 64));
  return v;
}

__device__ __forceinline__ uint32_t warp_inclusive_scan(uint32_t v) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
    if (lane >= d) v += t;
  }
  return v;
}

// emitted-row word: DROP keeps filter ∧ valid; EMIT_NULL keeps filter ∨ ¬valid
// (vector_selection.cc:61-88).  `inrange` masks rows beyond the array length.
__device__ __forceinline__ uint64_t emit_word64(uint64_t f, uint64_t fv, uint64_t inrange, int null_sel) {
  return null_sel ? ((f | ~fv) & inrange) : (f & fv);
}

__global__ void __launch_bounds__(kThreads)
filter_plan_kernel(BitmapRef f, BitmapRef fv, int64_t length, int null_sel, int64_t n_tiles, PlanView plan) {
  __shared__ uint32_t partial[16][8];
  __shared__ uint32_t tile_counts[kTilesPerSuper];
  __shared__ bool is_last;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int64_t super = blockIdx.x;
  const int64_t row0 = super * int64_t(kSuperRows);
  const BitmapRef ones{nullptr, 0};
  // word w = i*256 + t of the super-tile; tile of that word = i*4 + t/64
#pragma unroll 4
  for (int i = 0; i < 16; ++i) {
    int64_t pos = row0 + (int64_t(i) * 256 + t) * 64;
    uint32_t c = 0;
    if (pos < length) {
      uint64_t fw = bitmap_load64(f, pos, length);
      uint64_t vw = bitmap_load64(fv, pos, length);
      uint64_t in = bitmap_load64(ones, pos, length);
      c = __popcll(emit_word64(fw, vw, in, null_sel));
    }
    c = __reduce_add_sync(0xffffffffu, c);
    if (lane == 0) partial[i][warp] = c;
  }
  __syncthreads();
  uint32_t cnt = 0;
  if (t < kTilesPerSuper) {
    int i = t >> 2, q = t & 3;
    cnt = partial[i][2 * q] + partial[i][2 * q + 1];
    tile_counts[t] = cnt;
  }
  __syncthreads();
  if (t < kTilesPerSuper) {
    // exclusive scan over 64 counts: two warps
    uint32_t incl = warp_inclusive_scan(cnt);
    uint32_t add = 0;
    if (t >= 32) {
      for (int j = 0; j < 32; ++j) add += tile_counts[j];
    }
    int64_t tile = super * kTilesPerSuper + t;
    if (tile < n_tiles) plan.tile_prefix[tile] = incl - cnt + add;
    if (t == kTilesPerSuper - 1) plan.super_total[super] = int64_t(incl) + add;
  }
  // threadfence reduction: the last block scans the super-tile totals
  __threadfence();
  __syncthreads();
  if (t == 0) {
    uint32_t ticket = atomicAdd(&plan.header->done_counter, 1u);
    is_last = (ticket == gridDim.x - 1);
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  // single-block scan of n_super totals (<= a few thousand even at 1 B rows)
  __shared__ int64_t carry_s;
  __shared__ int64_t warp_sums[8];
  if (t == 0) carry_s = 0;
  __syncthreads();
  const int64_t ns = gridDim.x;
  const volatile int64_t* totals = plan.super_total;
  for (int64_t base = 0; base < ns; base += kThreads) {
    int64_t idx = base + t;
    int64_t v = idx < ns ? totals[idx] : 0;
    int64_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      int64_t o = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += o;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    int64_t wpre = 0;
    for (int w = 0; w < warp; ++w) wpre += warp_sums[w];
    int64_t carry = carry_s;
    if (idx < ns) plan.super_prefix[idx] = carry + wpre + incl - v;
    __syncthreads();
    if (t == kThreads - 1) carry_s = carry + wpre + incl;
    __syncthreads();
  }
  if (t == 0) {
    plan.header->out_length = carry_s;
    plan.header->done_counter = 0;
  }
}

struct ApplyArgs {
  BitmapRef f, fv, vv;   // mask data, mask validity, values validity
  BitmapRef vbits;       // values data as a bitmap (BOOL only)
  const void* vdata;     // values data, already shifted by the element offset (non-BOOL)
  int64_t length;
  int null_sel;
  const int64_t* super_prefix;
  const uint32_t* tile_prefix;
  void* out_data;
  uint32_t* out_valid;   // nullable
  const int64_t* out_length;   // the plan's number of emitted rows (device)
  int64_t capacity;            // rows the caller's output buffers hold
  uint32_t* err;               // context flag: set to 2 when the plan emits more rows than the buffers hold
};

// Appends compacted bit groups to an output bitmap as whole 32-bit words.  All lanes carry the
// same (warp-uniform) state; lane 0 performs the memory operations.  Words shared with a
// neighbouring warp/tile (first and last of the range) are merged with atomicOr into the
// zero-initialised output, interior words are plain stores.
struct BitAppender {
  uint32_t* out;
  int64_t word;
  uint64_t acc;
  int nacc;
  bool first_shared;
  __device__ __forceinline__ void init(uint32_t* o, int64_t start_bit) {
    out = o;
    word = start_bit >> 5;
    nacc = int(start_bit & 31);
    acc = 0;
    first_shared = nacc != 0;
  }
  __device__ __forceinline__ void append(uint32_t bits, int n, int lane) {
    acc |= uint64_t(bits) << nacc;
    nacc += n;
    if (nacc >= 32) {
      if (lane == 0) {
        uint32_t w = uint32_t(acc);
        if (first_shared) {
          if (w) atomicOr(out + word, w);
        } else {
          out[word] = w;
        }
      }
      first_shared = false;
      ++word;
      acc >>= 32;
      nacc -= 32;
    }
  }
  __device__ __forceinline__ void finish(int lane) {
    if (nacc > 0 && lane == 0) {
      uint32_t w = uint32_t(acc);
      if (w) atomicOr(out + word, w);
    }
  }
};

template <typename T, bool kBool, bool kHasValid>
__global__ void __launch_bounds__(kThreads)
filter_apply_kernel(ApplyArgs a) {
  __shared__ uint32_t warp_totals[8];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int64_t tile = blockIdx.x;
  // never write past the caller's buffers: an undersized output is reported through the context (b200_ctx_synchronize)
  if (*a.out_length > a.capacity) {
    if (tile == 0 && threadIdx.x == 0) atomicExch(a.err, 2u);
    return;
  }
  const int64_t out_base = a.super_prefix[tile / kTilesPerSuper] + a.tile_prefix[tile];
  const int64_t wrow0 = tile * int64_t(kTileRows) + int64_t(warp) * 512;
  const BitmapRef ones{nullptr, 0};

  uint32_t e = 0, iv = 0, vvw = 0, dbw = 0;
  if (lane < 16) {
    int64_t r = wrow0 + lane * 32;
    if (r < a.length) {
      uint32_t f = bitmap_load32(a.f, r, a.length);
      uint32_t fv = bitmap_load32(a.fv, r, a.length);
      uint32_t in = bitmap_load32(ones, r, a.length);
      iv = f & fv;
      e = a.null_sel ? ((f | ~fv) & in) : iv;
      if (kHasValid) vvw = bitmap_load32(a.vv, r, a.length);
      if (kBool) dbw = bitmap_load32(a.vbits, r, a.length);
    }
  }
  const uint32_t cnt = __popc(e);
  const uint32_t incl = warp_inclusive_scan(cnt);
  const uint32_t step_prefix = incl - cnt;
  const uint32_t wtotal = __shfl_sync(0xffffffffu, incl, 31);
  if (lane == 0) warp_totals[warp] = wtotal;
  __syncthreads();
  uint32_t wprefix = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w)
    if (w < warp) wprefix += warp_totals[w];
  if (wtotal == 0) return;  // warp-uniform: nothing emitted by this warp

  const int64_t wout = out_base + wprefix;
  BitAppender valid_app, data_app;
  if (kHasValid) valid_app.init(a.out_valid, wout);
  if (kBool) data_app.init(reinterpret_cast<uint32_t*>(a.out_data), wout);
  const T* vals = reinterpret_cast<const T*>(a.vdata);
  T* out = reinterpret_cast<T*>(a.out_data);
  const unsigned lt = lanemask_lt();

  // every value load of the warp's 512 rows is issued before the first store (16 independent
  // 256-byte requests in flight per warp); rows that are not emitted are not requested
  T vreg[16];
  if (!kBool) {
#pragma unroll
    for (int s = 0; s < 16; ++s) {
      const uint32_t ivs = __shfl_sync(0xffffffffu, iv, s);
      vreg[s] = T(0);  // an emitted filter-null writes zero (vector_selection.cc:743-746)
      if ((ivs >> lane) & 1u) vreg[s] = __ldg(vals + wrow0 + s * 32 + lane);
    }
  }
#pragma unroll
  for (int s = 0; s < 16; ++s) {
    const uint32_t es = __shfl_sync(0xffffffffu, e, s);
    if (es == 0) continue;  // warp-uniform
    const uint32_t ivs = __shfl_sync(0xffffffffu, iv, s);
    const uint32_t ps = __shfl_sync(0xffffffffu, step_prefix, s);
    const bool emit = (es >> lane) & 1u;
    const bool isv = (ivs >> lane) & 1u;
    const uint32_t rank = __popc(es & lt);
    if (!kBool) {
      if (emit) out[wout + ps + rank] = vreg[s];
    } else {
      const uint32_t ds = __shfl_sync(0xffffffffu, dbw, s);
      const uint32_t bit = (isv && ((ds >> lane) & 1u)) ? (1u << rank) : 0u;
      const uint32_t cw = __reduce_or_sync(0xffffffffu, bit);
      data_app.append(cw, __popc(es), lane);
    }
    if (kHasValid) {
      const uint32_t vs = __shfl_sync(0xffffffffu, vvw, s);
      const uint32_t bit = (isv && ((vs >> lane) & 1u)) ? (1u << rank) : 0u;
      const uint32_t cw = __reduce_or_sync(0xffffffffu, bit);
      valid_app.append(cw, __popc(es), lane);
    }
  }
  if (kHasValid) valid_app.finish(lane);
  if (kBool) data_app.finish(lane);
}


// Measured and dropped in round 2 (profiles/r02_filter_variants.txt):
//  * one launch with the tile offsets taken by decoupled look-back instead of from filter_plan_kernel's prefix arrays:
//    0.41 ms per 100 M rows against 0.02 + 0.25 ms — with ~1200 short tiles in flight a new tile walks dozens of
//    look-back windows before it meets an inclusive state, while the plan kernel only reads the 12.5 MB mask;
//  * two rows per lane with unconditional LDG.E.128 value loads: 0.30 ms against 0.25 ms (every sector is fetched
//    either way at 50 % selectivity, and each lane then issues two scattered 8-byte stores per load).
// b200_filter_fused therefore is filter_plan + filter_apply back to back on the stream, without the host round trip.

int check_mask(b200_ctx* ctx, const b200_col* mask) {
  if (!mask) return b200_fail(ctx, B200_INVALID, "filter: null mask descriptor");
  // FilterMetaFunction: "Filter argument must be boolean type" (vector_selection.cc:1885-1887)
  if (mask->type != B200_BOOL)
    return b200_fail(ctx, B200_NOT_IMPLEMENTED, "Filter argument must be boolean type");
  if (mask->length < 0 || mask->offset < 0) return b200_fail(ctx, B200_INVALID, "filter: negative length/offset");
  if (mask->length > 0 && mask->data == nullptr) return b200_fail(ctx, B200_INVALID, "filter: mask has no data buffer");
  return B200_OK;
}

template <typename T, bool kBool>
int launch_apply(b200_ctx* ctx, const ApplyArgs& a, int64_t n_tiles) {
  dim3 grid((unsigned)n_tiles), block(kThreads);
  if (a.out_valid)
    B200_TIMED(ctx, "filter_apply", (filter_apply_kernel<T, kBool, true><<<grid, block, 0, ctx->stream>>>(a)));
  else
    B200_TIMED(ctx, "filter_apply", (filter_apply_kernel<T, kBool, false><<<grid, block, 0, ctx->stream>>>(a)));
  return B200_OK;
}

// ---- mask producer: values[i] OP scalar -> boolean bitmap (compute/kernels/scalar_compare.cc) -----------
// One row per thread; the warp's 32 predicate bits are one ballot = one bitmap word (no atomics, no
// shared memory).  Arrow semantics: C++ comparison operators on the physical values (a NaN compares
// false except under not_equal); null-ness is not computed here: the output shares the input's validity.
enum { kCmpEq = 0, kCmpNe = 1, kCmpGt = 2, kCmpGe = 3, kCmpLt = 4, kCmpLe = 5 };

template <typename T>
__device__ __forceinline__ bool cmp_apply(T v, T s, int op) {
  switch (op) {
    case kCmpEq: return v == s;
    case kCmpNe: return v != s;
    case kCmpGt: return v > s;
    case kCmpGe: return v >= s;
    case kCmpLt: return v < s;
    default: return v <= s;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) compare_scalar_kernel(const T* __restrict__ values, int64_t n, T scalar, int op,
                                                            uint32_t* __restrict__ out) {
  // 4 words (128 rows) per warp iteration keeps enough loads in flight for narrow types
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t base = warp * 128;
  if (base >= n) return;
  T v[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int64_t row = base + k * 32 + lane;
    v[k] = row < n ? values[row] : T(0);
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int64_t row = base + k * 32 + lane;
    const unsigned word = __ballot_sync(0xffffffffu, row < n && cmp_apply<T>(v[k], scalar, op));
    if (lane == 0 && base + k * 32 < n) out[(base >> 5) + k] = word;
  }
}

template <typename T>
int launch_compare(b200_ctx* ctx, const b200_col* values, const void* scalar, int op, void* out) {
  T s;
  memcpy(&s, scalar, sizeof(T));
  const int64_t n = values->length;
  const T* data = reinterpret_cast<const T*>(values->data) + values->offset;
  const int64_t warps = (n + 127) / 128;
  const unsigned grid = unsigned((warps * 32 + 255) / 256);
  B200_TIMED(ctx, "compare_scalar", (compare_scalar_kernel<T><<<grid, 256, 0, ctx->stream>>>(data, n, s, op, reinterpret_cast<uint32_t*>(out))));
  return B200_OK;
}

// ---- Take: out[i] = values[indices[i]] (vector_selection.cc:273-469)
struct TakeArgs {
  const void* vdata;       // values, shifted by the element offset (not for bool)
  BitmapRef vbits;         // bool values
  BitmapRef vv;            // values validity
  int64_t vlen;
  const void* idata;       // indices, shifted by the element offset
  BitmapRef iv;            // indices validity
  int32_t itype;
  int64_t n;
  void* out;
  uint32_t* out_valid;
  unsigned long long* counters;   // [0] valid rows, [1] first row with an out-of-bounds index (or ~0)
  // chunked values (b200_take_chunked): chunk c holds logical rows [starts[c], starts[c + 1])
  const struct TakeChunk* chunks;
  const int64_t* starts;          // device, n_chunks + 1 entries
  int32_t n_chunks;
};

struct TakeChunk {
  const void* vdata;   // shifted by the chunk's element offset (not for bool)
  BitmapRef vbits;     // bool values
  BitmapRef vv;        // validity (words == nullptr: no nulls)
};

constexpr int kTakeSharedStarts = 1024;   // chunk boundaries kept in shared memory up to this many chunks

// the index as an unsigned number, and whether it is non-negative (I is the index C type: no per-row dispatch,
// so the compiler can batch the loads of a thread's rows)
template <typename I>
__device__ __forceinline__ bool take_index(const TakeArgs& a, int64_t row, uint64_t* ix) {
  const I v = reinterpret_cast<const I*>(a.idata)[row];
  *ix = std::is_signed<I>::value ? uint64_t(int64_t(v)) : uint64_t(v);
  return !std::is_signed<I>::value || v >= I(0);
}

constexpr int kTakeItems = 8;   // gathers in flight per thread

template <typename V, bool kBool, typename I, bool kChunked = false>
__global__ void __launch_bounds__(256) take_kernel(const __grid_constant__ TakeArgs a) {
  // chunked values: the boundaries go to shared memory once per CTA (global memory beyond 1024 chunks)
  __shared__ int64_t s_starts[kChunked ? kTakeSharedStarts + 1 : 1];
  const int64_t* starts = a.starts;
  if (kChunked && a.n_chunks <= kTakeSharedStarts) {
    for (int c = threadIdx.x; c <= a.n_chunks; c += blockDim.x) s_starts[c] = a.starts[c];
    __syncthreads();
    starts = s_starts;
  }
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t base = warp * (32 * kTakeItems);
  if (base >= a.n) return;
  // three batches of independent loads, no branch between the loads of a batch: index validity words and
  // indices; then value validity bytes and values (addresses clamped to row 0 where the index is unusable)
  uint64_t ix[kTakeItems];
  bool ok[kTakeItems];
  uint32_t ivw[kTakeItems];
#pragma unroll
  for (int k = 0; k < kTakeItems; ++k) ivw[k] = bitmap_load32(a.iv, base + k * 32, a.n);
  bool nonneg[kTakeItems];
#pragma unroll
  for (int k = 0; k < kTakeItems; ++k) {
    const int64_t row = min(base + k * 32 + lane, a.n - 1);
    nonneg[k] = take_index<I>(a, row, &ix[k]);
  }
#pragma unroll
  for (int k = 0; k < kTakeItems; ++k) {
    const int64_t row = base + k * 32 + lane;
    const bool present = row < a.n && ((ivw[k] >> lane) & 1u);
    ok[k] = present && nonneg[k] && ix[k] < uint64_t(a.vlen);
    if (present && !ok[k]) atomicMin(a.counters + 1, (unsigned long long)row);
    if (!ok[k]) ix[k] = 0;
  }
  V val[kTakeItems];
  bool vvalid[kTakeItems];
  if (kChunked) {
    // the last chunk whose first row is <= the index (empty chunks share a boundary with their successor and
    // are never picked); the eight searches of a thread are independent and interleave
    int chunk[kTakeItems];
#pragma unroll
    for (int k = 0; k < kTakeItems; ++k) {
      int lo = 0, hi = a.n_chunks;          // invariant: starts[lo] <= ix < starts[hi]
      while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (uint64_t(starts[mid]) <= ix[k]) lo = mid; else hi = mid;
      }
      chunk[k] = lo;
      ix[k] -= uint64_t(starts[lo]);
    }
#pragma unroll
    for (int k = 0; k < kTakeItems; ++k) {
      const TakeChunk c = a.chunks[chunk[k]];
      vvalid[k] = bitmap_get(c.vv, int64_t(ix[k]));
      val[k] = kBool ? V(bitmap_get(c.vbits, int64_t(ix[k]))) : reinterpret_cast<const V*>(c.vdata)[ix[k]];
    }
  } else {
#pragma unroll
    for (int k = 0; k < kTakeItems; ++k) {
      vvalid[k] = bitmap_get(a.vv, int64_t(ix[k]));
      val[k] = kBool ? V(bitmap_get(a.vbits, int64_t(ix[k]))) : reinterpret_cast<const V*>(a.vdata)[ix[k]];
    }
  }
#pragma unroll
  for (int k = 0; k < kTakeItems; ++k) {
    ok[k] = ok[k] && vvalid[k];
    if (!ok[k]) val[k] = V(0);
  }
  uint32_t valid_rows = 0;
#pragma unroll
  for (int k = 0; k < kTakeItems; ++k) {
    const int64_t row = base + k * 32 + lane;
    const unsigned word = __ballot_sync(0xffffffffu, ok[k]);
    if (kBool) {
      const unsigned bits = __ballot_sync(0xffffffffu, ok[k] && val[k] != V(0));
      if (lane == 0 && base + k * 32 < a.n) reinterpret_cast<uint32_t*>(a.out)[(base >> 5) + k] = bits;
    } else if (row < a.n) {
      reinterpret_cast<V*>(a.out)[row] = val[k];
    }
    if (lane == 0 && base + k * 32 < a.n) a.out_valid[(base >> 5) + k] = word;
    valid_rows += __popc(word);
  }
  if (lane == 0 && valid_rows) atomicAdd(a.counters, (unsigned long long)valid_rows);
}

template <typename V, bool kBool, typename I>
int launch_take_i(b200_ctx* ctx, const TakeArgs& a) {
  const int64_t warps = (a.n + 32 * kTakeItems - 1) / (32 * kTakeItems);
  const unsigned grid = unsigned((warps * 32 + 255) / 256);
  if (a.chunks) B200_TIMED(ctx, "take", (take_kernel<V, kBool, I, true><<<grid, 256, 0, ctx->stream>>>(a)));
  else B200_TIMED(ctx, "take", (take_kernel<V, kBool, I><<<grid, 256, 0, ctx->stream>>>(a)));
  return B200_OK;
}

template <typename V, bool kBool>
int launch_take(b200_ctx* ctx, const TakeArgs& a) {
  switch (a.itype) {
    case B200_INT8: return launch_take_i<V, kBool, int8_t>(ctx, a);
    case B200_UINT8: return launch_take_i<V, kBool, uint8_t>(ctx, a);
    case B200_INT16: return launch_take_i<V, kBool, int16_t>(ctx, a);
    case B200_UINT16: return launch_take_i<V, kBool, uint16_t>(ctx, a);
    case B200_INT32: return launch_take_i<V, kBool, int32_t>(ctx, a);
    case B200_UINT32: return launch_take_i<V, kBool, uint32_t>(ctx, a);
    case B200_INT64: return launch_take_i<V, kBool, int64_t>(ctx, a);
    default: return launch_take_i<V, kBool, uint64_t>(ctx, a);
  }
}

// ---- add / subtract / multiply (scalar_arithmetic.cc, unchecked forms)
template <typename T>
struct ArithTraits {
  // integers: computed in the unsigned type of the same width so that overflow wraps (and 16-bit products do
  // not promote to a signed int that could overflow), like the reference's SafeSigned* / to_unsigned forms
  using U = typename std::make_unsigned<T>::type;
  static __device__ __forceinline__ T apply(int op, T x, T y) {
    const U ux = U(x), uy = U(y);
    return T(op == 0 ? U(ux + uy) : (op == 1 ? U(ux - uy) : U(ux * uy)));
  }
};
template <>
struct ArithTraits<float> {
  static __device__ __forceinline__ float apply(int op, float x, float y) {
    return op == 0 ? __fadd_rn(x, y) : (op == 1 ? __fsub_rn(x, y) : __fmul_rn(x, y));   // never contracted
  }
};
template <>
struct ArithTraits<double> {
  static __device__ __forceinline__ double apply(int op, double x, double y) {
    return op == 0 ? __dadd_rn(x, y) : (op == 1 ? __dsub_rn(x, y) : __dmul_rn(x, y));
  }
};

constexpr int kArithItems = 4;

template <typename T>
__global__ void __launch_bounds__(256) arith_kernel(const T* __restrict__ a, const T* __restrict__ b, T sa, T sb, int op,
                                                    int64_t n, T* __restrict__ out) {
  const int64_t base = (int64_t(blockIdx.x) * 256) * kArithItems + threadIdx.x;
  T x[kArithItems], y[kArithItems];
#pragma unroll
  for (int k = 0; k < kArithItems; ++k) {
    const int64_t i = base + k * 256;
    x[k] = (a && i < n) ? a[i] : sa;
    y[k] = (b && i < n) ? b[i] : sb;
  }
#pragma unroll
  for (int k = 0; k < kArithItems; ++k) {
    const int64_t i = base + k * 256;
    if (i < n) out[i] = ArithTraits<T>::apply(op, x[k], y[k]);
  }
}

// out word w = validity bits [32w, 32w+32) of a AND of b (a missing bitmap reads as all ones)
__global__ void __launch_bounds__(256) bitmap_and_kernel(BitmapRef a, BitmapRef b, int64_t n, uint32_t* __restrict__ out) {
  const int64_t w = int64_t(blockIdx.x) * 256 + threadIdx.x;
  if (w * 32 < n) out[w] = bitmap_load32(a, w * 32, n) & bitmap_load32(b, w * 32, n);
}

template <typename T>
int launch_arith(b200_ctx* ctx, int op, const b200_col* a, const void* a_scalar, const b200_col* b, const void* b_scalar,
                 int64_t n, void* out) {
  T sa = T(0), sb = T(0);
  if (!a) memcpy(&sa, a_scalar, sizeof(T));
  if (!b) memcpy(&sb, b_scalar, sizeof(T));
  const T* pa = a ? reinterpret_cast<const T*>(a->data) + a->offset : nullptr;
  const T* pb = b ? reinterpret_cast<const T*>(b->data) + b->offset : nullptr;
  const unsigned grid = unsigned((n + 256 * kArithItems - 1) / (256 * kArithItems));
  B200_TIMED(ctx, "arithmetic", (arith_kernel<T><<<grid, 256, 0, ctx->stream>>>(pa, pb, sa, sb, op, n, reinterpret_cast<T*>(out))));
  return B200_OK;
}

}  // namespace

extern "C" {

// values: one column (n_chunks == 0) or the chunks of one (n_chunks >= 1, all of one type)
static int take_impl(b200_ctx* ctx, const b200_col* values, int n_chunks, const b200_col* indices, int boundscheck,
                     void* out_data, void* out_validity, int64_t* out_null_count) {
  const int vbits = b200_type_bits(values->type);
  if (vbits == 0) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "take: unsupported values type %d", values->type);
  const int ibits = b200_type_bits(indices->type);
  if (ibits < 8 || indices->type == B200_FLOAT || indices->type == B200_DOUBLE)
    return b200_fail(ctx, B200_TYPE_ERROR, "take: indices must be an integer column, got type %d", indices->type);
  if (indices->length < 0 || indices->offset < 0) return b200_fail(ctx, B200_INVALID, "take: negative length/offset");
  const int n_cols = n_chunks > 0 ? n_chunks : 1;
  int64_t vlen = 0;
  for (int c = 0; c < n_cols; ++c) {
    if (values[c].type != values->type) return b200_fail(ctx, B200_TYPE_ERROR, "take: chunks of different types");
    if (values[c].length < 0 || values[c].offset < 0) return b200_fail(ctx, B200_INVALID, "take: negative length/offset");
    if (values[c].length > 0 && !values[c].data) return b200_fail(ctx, B200_INVALID, "take: null values buffer");
    vlen += values[c].length;
  }
  const int64_t n = indices->length;
  if (out_null_count) *out_null_count = 0;
  if (n == 0) return B200_OK;
  if (!indices->data || !out_data || !out_validity)
    return b200_fail(ctx, B200_INVALID, "take: null buffer (the output validity bitmap is always produced)");
  const size_t starts_bytes = (size_t(n_cols) + 1) * sizeof(int64_t);
  B200_RETURN_NOT_OK(b200_ctx_reserve_scratch(ctx, 256 + starts_bytes + size_t(n_cols) * sizeof(TakeChunk)));
  unsigned long long* counters = reinterpret_cast<unsigned long long*>(ctx->scratch);
  const unsigned long long init[2] = {0ull, ~0ull};
  ctx->host_words[8] = int64_t(init[0]);
  ctx->host_words[9] = int64_t(init[1]);
  B200_CUDA(ctx, cudaMemcpyAsync(counters, ctx->host_words + 8, 16, cudaMemcpyHostToDevice, ctx->stream));
  TakeArgs a;
  a.chunks = nullptr;
  a.starts = nullptr;
  a.n_chunks = 0;
  a.vlen = vlen;
  if (vlen == 0) {   // every index is out of range; the clamped gathers read row 0 of the scratch block
    a.vdata = counters;
    a.vbits = make_bitmap_ref(counters, 0);
    a.vv = BitmapRef{nullptr, 0};
  } else if (n_chunks <= 1) {
    a.vdata = vbits >= 8 ? reinterpret_cast<const char*>(values->data) + values->offset * (vbits / 8) : nullptr;
    a.vbits = make_bitmap_ref(values->data, values->offset);
    a.vv = (values->validity && values->null_count != 0) ? make_bitmap_ref(values->validity, values->offset) : BitmapRef{nullptr, 0};
  } else {
    // chunk table: pageable source, so the runtime has staged it by the time the copy call returns
    std::vector<int64_t> starts(size_t(n_cols) + 1);
    std::vector<TakeChunk> table(n_cols);
    int64_t at = 0;
    for (int c = 0; c < n_cols; ++c) {
      const b200_col& v = values[c];
      starts[c] = at;
      at += v.length;
      table[c].vdata = vbits >= 8 ? reinterpret_cast<const char*>(v.data) + v.offset * (vbits / 8) : nullptr;
      table[c].vbits = make_bitmap_ref(v.data, v.offset);
      table[c].vv = (v.validity && v.null_count != 0) ? make_bitmap_ref(v.validity, v.offset) : BitmapRef{nullptr, 0};
    }
    starts[n_cols] = at;
    char* base = reinterpret_cast<char*>(ctx->scratch) + 256;
    B200_CUDA(ctx, cudaMemcpyAsync(base, starts.data(), starts_bytes, cudaMemcpyHostToDevice, ctx->stream));
    B200_CUDA(ctx, cudaMemcpyAsync(base + starts_bytes, table.data(), table.size() * sizeof(TakeChunk), cudaMemcpyHostToDevice, ctx->stream));
    a.starts = reinterpret_cast<const int64_t*>(base);
    a.chunks = reinterpret_cast<const TakeChunk*>(base + starts_bytes);
    a.n_chunks = n_cols;
    a.vdata = nullptr;
    a.vbits = BitmapRef{nullptr, 0};
    a.vv = BitmapRef{nullptr, 0};
  }
  a.idata = reinterpret_cast<const char*>(indices->data) + indices->offset * (ibits / 8);
  a.iv = (indices->validity && indices->null_count != 0) ? make_bitmap_ref(indices->validity, indices->offset) : BitmapRef{nullptr, 0};
  a.itype = indices->type;
  a.n = n;
  a.out = out_data;
  a.out_valid = reinterpret_cast<uint32_t*>(out_validity);
  a.counters = counters;
  switch (vbits) {
    case 1: B200_RETURN_NOT_OK((launch_take<uint8_t, true>(ctx, a))); break;
    case 8: B200_RETURN_NOT_OK((launch_take<uint8_t, false>(ctx, a))); break;
    case 16: B200_RETURN_NOT_OK((launch_take<uint16_t, false>(ctx, a))); break;
    case 32: B200_RETURN_NOT_OK((launch_take<uint32_t, false>(ctx, a))); break;
    default: B200_RETURN_NOT_OK((launch_take<uint64_t, false>(ctx, a))); break;
  }
  if (boundscheck || out_null_count) {
    B200_CUDA(ctx, cudaMemcpyAsync(ctx->host_words + 8, counters, 16, cudaMemcpyDeviceToHost, ctx->stream));
    B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const uint64_t bad_row = uint64_t(ctx->host_words[9]);
    if (boundscheck && bad_row != ~uint64_t(0)) {
      // the offending value, for the reference's message (error path only)
      int64_t raw = 0;
      B200_CUDA(ctx, cudaMemcpy(&raw, reinterpret_cast<const char*>(a.idata) + bad_row * (ibits / 8), ibits / 8, cudaMemcpyDeviceToHost));
      long long shown;
      switch (indices->type) {
        case B200_INT8: shown = (long long)(int8_t)raw; break;
        case B200_INT16: shown = (long long)(int16_t)raw; break;
        case B200_INT32: shown = (long long)(int32_t)raw; break;
        default: shown = (long long)raw; break;   // little-endian: narrower unsigned values are zero-extended
      }
      if (indices->type == B200_UINT64) return b200_fail(ctx, B200_INDEX_ERROR, "Index %llu out of bounds", (unsigned long long)raw);
      return b200_fail(ctx, B200_INDEX_ERROR, "Index %lld out of bounds", shown);
    }
    if (out_null_count) *out_null_count = n - ctx->host_words[8];
  }
  return B200_OK;
}

int b200_take(b200_ctx* ctx, const b200_col* values, const b200_col* indices, int boundscheck, void* out_data,
              void* out_validity, int64_t* out_null_count) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (!values || !indices) return b200_fail(ctx, B200_INVALID, "take: null argument");
  return take_impl(ctx, values, 0, indices, boundscheck, out_data, out_validity, out_null_count);
}

int b200_take_chunked(b200_ctx* ctx, const b200_col* chunks, int n_chunks, const b200_col* indices, int boundscheck,
                      void* out_data, void* out_validity, int64_t* out_null_count) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (!chunks || !indices || n_chunks < 1) return b200_fail(ctx, B200_INVALID, "take_chunked: null argument or no chunks");
  return take_impl(ctx, chunks, n_chunks, indices, boundscheck, out_data, out_validity, out_null_count);
}

int b200_arithmetic(b200_ctx* ctx, int op, const b200_col* a, const void* a_scalar, const b200_col* b, const void* b_scalar,
                    void* out_data, void* out_validity) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (op < 0 || op > 2) return b200_fail(ctx, B200_INVALID, "arithmetic: bad operator %d", op);
  if (!a && !b) return b200_fail(ctx, B200_INVALID, "arithmetic: at least one operand must be a column");
  if ((!a && !a_scalar) || (!b && !b_scalar)) return b200_fail(ctx, B200_INVALID, "arithmetic: missing operand");
  const b200_col* c = a ? a : b;
  if (a && b) {
    // ExecBatchIterator: "Array arguments must all be the same length" (compute/exec.cc:186-189)
    if (a->length != b->length) return b200_fail(ctx, B200_INVALID, "Array arguments must all be the same length");
    // the reference casts to a common type first (DispatchBest); that cast is not part of this path
    if (a->type != b->type) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "arithmetic: operand types differ (%d, %d)", a->type, b->type);
  }
  const int64_t n = c->length;
  if (n < 0 || c->offset < 0 || (a && b && b->offset < 0)) return b200_fail(ctx, B200_INVALID, "arithmetic: negative length/offset");
  if (n == 0) return B200_OK;
  if (!c->data || (a && b && !b->data) || !out_data) return b200_fail(ctx, B200_INVALID, "arithmetic: null buffer");
  if (out_validity) {
    const BitmapRef va = a ? make_bitmap_ref(a->validity, a->offset) : BitmapRef{nullptr, 0};
    const BitmapRef vb = b ? make_bitmap_ref(b->validity, b->offset) : BitmapRef{nullptr, 0};
    const int64_t words = (n + 31) / 32;
    bitmap_and_kernel<<<unsigned((words + 255) / 256), 256, 0, ctx->stream>>>(va, vb, n, reinterpret_cast<uint32_t*>(out_validity));
    B200_LAUNCHED(ctx);
  }
  switch (c->type) {
    case B200_UINT8: return launch_arith<uint8_t>(ctx, op, a, a_scalar, b, b_scalar, n, out_data);
    case B200_INT8: return launch_arith<int8_t>(ctx, op, a, a_scalar, b, b_scalar, n, out_data);
    case B200_UINT16: return launch_arith<uint16_t>(ctx, op, a, a_scalar, b, b_scalar, n, out_data);
    case B200_INT16: return launch_arith<int16_t>(ctx, op, a, a_scalar, b, b_scalar, n, out_data);
    case B200_UINT32: return launch_arith<uint32_t>(ctx, op, a, a_scalar, b, b_scalar, n, out_data);
    case B200_INT32: return launch_arith<int32_t>(ctx, op, a, a_scalar, b, b_scalar, n, out_data);
    case B200_UINT64: return launch_arith<uint64_t>(ctx, op, a, a_scalar, b, b_scalar, n, out_data);
    case B200_INT64: return launch_arith<int64_t>(ctx, op, a, a_scalar, b, b_scalar, n, out_data);
    case B200_FLOAT: return launch_arith<float>(ctx, op, a, a_scalar, b, b_scalar, n, out_data);
    case B200_DOUBLE: return launch_arith<double>(ctx, op, a, a_scalar, b, b_scalar, n, out_data);
    default: break;
  }
  return b200_fail(ctx, B200_NOT_IMPLEMENTED, "arithmetic: no kernel for type %d", c->type);
}

int b200_filter_plan_bytes(int64_t mask_length, size_t* out_bytes) {
  if (!out_bytes || mask_length < 0) return B200_INVALID;
  *out_bytes = plan_bytes_for(mask_length);
  return B200_OK;
}

const int64_t* b200_filter_plan_length_ptr(const void* plan) {
  return &reinterpret_cast<const PlanHeader*>(plan)->out_length;
}

int b200_filter_plan(b200_ctx* ctx, const b200_col* mask, int null_selection, void* plan, int64_t* out_len) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  B200_RETURN_NOT_OK(check_mask(ctx, mask));
  if (!plan) return b200_fail(ctx, B200_INVALID, "filter_plan: null plan buffer");
  if (null_selection != 0 && null_selection != 1) return b200_fail(ctx, B200_INVALID, "filter: bad null_selection");
  const int64_t length = mask->length;
  PlanView pv = plan_view(plan, length);
  // out_length = 0 and done_counter = 0 before the scan starts
  B200_CUDA(ctx, cudaMemsetAsync(pv.header, 0, sizeof(PlanHeader), ctx->stream));
  if (length > 0) {
    BitmapRef f = make_bitmap_ref(mask->data, mask->offset);
    BitmapRef fv = make_bitmap_ref(mask->validity, mask->offset);
    B200_TIMED(ctx, "filter_plan",
               (filter_plan_kernel<<<(unsigned)n_super_for(length), kThreads, 0, ctx->stream>>>(
                   f, fv, length, null_selection, n_tiles_for(length), pv)));
  }
  if (out_len) {
    B200_CUDA(ctx, cudaMemcpyAsync(ctx->host_words, &pv.header->out_length, sizeof(int64_t),
                                   cudaMemcpyDeviceToHost, ctx->stream));
    B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *out_len = ctx->host_words[0];
  }
  return B200_OK;
}

int b200_compare_scalar(b200_ctx* ctx, const b200_col* values, int op, const void* scalar, void* out_bitmap) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (!values || !scalar) return b200_fail(ctx, B200_INVALID, "compare_scalar: null argument");
  if (op < kCmpEq || op > kCmpLe) return b200_fail(ctx, B200_INVALID, "compare_scalar: bad operator %d", op);
  if (values->length < 0 || values->offset < 0) return b200_fail(ctx, B200_INVALID, "compare_scalar: negative length/offset");
  if (values->length == 0) return B200_OK;
  if (!values->data || !out_bitmap) return b200_fail(ctx, B200_INVALID, "compare_scalar: null buffer");
  switch (values->type) {
    case B200_UINT8: return launch_compare<uint8_t>(ctx, values, scalar, op, out_bitmap);
    case B200_INT8: return launch_compare<int8_t>(ctx, values, scalar, op, out_bitmap);
    case B200_UINT16: return launch_compare<uint16_t>(ctx, values, scalar, op, out_bitmap);
    case B200_INT16: return launch_compare<int16_t>(ctx, values, scalar, op, out_bitmap);
    case B200_UINT32: return launch_compare<uint32_t>(ctx, values, scalar, op, out_bitmap);
    case B200_INT32: return launch_compare<int32_t>(ctx, values, scalar, op, out_bitmap);
    case B200_UINT64: return launch_compare<uint64_t>(ctx, values, scalar, op, out_bitmap);
    case B200_INT64: return launch_compare<int64_t>(ctx, values, scalar, op, out_bitmap);
    case B200_FLOAT: return launch_compare<float>(ctx, values, scalar, op, out_bitmap);
    case B200_DOUBLE: return launch_compare<double>(ctx, values, scalar, op, out_bitmap);
    default: break;
  }
  // the reference has boolean comparisons too (scalar_compare.cc); not on the Q1 path
  return b200_fail(ctx, B200_NOT_IMPLEMENTED, "compare_scalar: no kernel for type %d", values->type);
}

int b200_filter_count(b200_ctx* ctx, const b200_col* mask, int null_selection, int64_t* out_len) {
  B200_ENTER(ctx);
  if (!ctx || !out_len) return B200_INVALID;
  B200_RETURN_NOT_OK(check_mask(ctx, mask));
  size_t bytes = plan_bytes_for(mask->length);
  B200_RETURN_NOT_OK(b200_ctx_reserve_scratch(ctx, bytes));
  return b200_filter_plan(ctx, mask, null_selection, ctx->scratch, out_len);
}

int b200_filter_fused_scratch_bytes(int64_t mask_length, size_t* out_bytes) { return b200_filter_plan_bytes(mask_length, out_bytes); }

int b200_filter_fused(b200_ctx* ctx, const b200_col* values, const b200_col* mask, int null_selection, void* scratch,
                      void* out_data, void* out_validity, int64_t out_capacity, int64_t* out_len) {
  // the plan (a scan of the mask alone: 1/64 of the value bytes) and the compaction, back to back on the stream;
  // the output length only travels to the host when the caller asks for it
  B200_RETURN_NOT_OK(b200_filter_plan(ctx, mask, null_selection, scratch, nullptr));
  B200_RETURN_NOT_OK(b200_filter_apply(ctx, values, mask, null_selection, scratch, out_data, out_validity, out_capacity));
  if (out_len) {
    B200_ENTER(ctx);
    *out_len = 0;
    if (mask->length == 0) return B200_OK;
    B200_CUDA(ctx, cudaMemcpyAsync(ctx->host_words, b200_filter_plan_length_ptr(scratch), sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *out_len = ctx->host_words[0];
  }
  return B200_OK;
}

// ---- Filter for callers that must not allocate --------------------------------------------------------------------------
// b200_filter_fused with its plan in a file-scope device array instead of caller scratch: no cudaMalloc, no cudaFree, no
// allocating library call anywhere on the path (CUDA graph capture, serving loops, harnesses that forbid allocation).
// The array costs kStaticPlanBytes on every GPU that loads the library, until the process exits.
constexpr int64_t kStaticMaxRows = int64_t(1) << 28;   // 268 435 456 mask rows
constexpr size_t kStaticPlanBytes = sizeof(PlanHeader) + size_t(kStaticMaxRows / kSuperRows) * 16 + size_t(kStaticMaxRows / kTileRows) * 4 + 64 + 256;
static __device__ __align__(256) unsigned char g_static_plan[kStaticPlanBytes];   // 272.4 KB

// In host code `g_static_plan` is only a stand-in for the symbol, not a device address: the address is asked for once
// per device, with that device current (B200_ENTER has made it so), and cached.
static int static_plan_address(b200_ctx* ctx, void** out) {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static void* addr[kMaxDevices] = {};
  if (ctx->device < 0 || ctx->device >= kMaxDevices) return b200_fail(ctx, B200_INVALID, "filter_fused_noalloc: device ordinal %d", ctx->device);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[ctx->device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_plan);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return b200_fail(ctx, B200_UNKNOWN_ERROR, "filter_fused_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
    }
    addr[ctx->device] = p;
  }
  *out = addr[ctx->device];
  return B200_OK;
}

int64_t b200_filter_fused_noalloc_max_rows(void) { return kStaticMaxRows; }

int b200_filter_fused_noalloc(b200_ctx* ctx, const b200_col* values, const b200_col* mask, int null_selection, void* out_data,
                              void* out_validity, int64_t out_capacity, int64_t* out_len) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  B200_RETURN_NOT_OK(check_mask(ctx, mask));
  if (mask->length > kStaticMaxRows || plan_bytes_for(mask->length) > kStaticPlanBytes)
    return b200_fail(ctx, B200_CAPACITY_ERROR, "filter_fused_noalloc: %lld rows exceed the built-in plan's maximum of %lld (use b200_filter_fused with caller scratch)",
                     (long long)mask->length, (long long)kStaticMaxRows);
  void* plan = nullptr;
  B200_RETURN_NOT_OK(static_plan_address(ctx, &plan));
  // the array keeps the previous call's contents: b200_filter_plan zeroes the header (output length, done counter) on
  // the stream and rewrites every prefix entry the apply kernel reads
  return b200_filter_fused(ctx, values, mask, null_selection, plan, out_data, out_validity, out_capacity, out_len);
}

int b200_filter_out_spec(const b200_col* values, const b200_col* mask, int null_selection,
                         int* allocate_validity, int64_t* out_null_count) {
  if (!values || !mask) return B200_INVALID;
  // GetPrimitiveArg: null_count is 0 when there is no validity bitmap (util_internal.cc:56)
  int64_t vnc = values->validity ? values->null_count : 0;
  int64_t fnc = mask->validity ? mask->null_count : 0;
  // PrimitiveFilter (vector_selection.cc:799-812)
  if (out_null_count) *out_null_count = (vnc == 0 && (null_selection == 0 || fnc == 0)) ? 0 : -1;
  if (allocate_validity) *allocate_validity = (vnc != 0 || fnc != 0) ? 1 : 0;
  return B200_OK;
}

int b200_filter_apply(b200_ctx* ctx, const b200_col* values, const b200_col* mask, int null_selection,
                      const void* plan, void* out_data, void* out_validity, int64_t out_capacity) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  B200_RETURN_NOT_OK(check_mask(ctx, mask));
  if (!values || !plan) return b200_fail(ctx, B200_INVALID, "filter_apply: null argument");
  // ExecBatchIterator: "Array arguments must all be the same length" (compute/exec.cc:186-189)
  if (values->length != mask->length)
    return b200_fail(ctx, B200_INVALID, "Array arguments must all be the same length");
  const int bits = b200_type_bits(values->type);
  if (bits == 0) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "filter: unsupported values type %d", values->type);
  if (out_capacity < 0) return b200_fail(ctx, B200_INVALID, "filter_apply: negative out_capacity");
  const int64_t length = mask->length;
  if (length == 0) return B200_OK;
  if (!out_data && out_capacity > 0) return b200_fail(ctx, B200_INVALID, "filter_apply: null out_data");
  if ((reinterpret_cast<uintptr_t>(out_data) & 7) || (reinterpret_cast<uintptr_t>(out_validity) & 7))
    return b200_fail(ctx, B200_INVALID, "filter_apply: output buffers must be 8-byte aligned");
  PlanView pv = plan_view(const_cast<void*>(plan), length);
  ApplyArgs a;
  a.f = make_bitmap_ref(mask->data, mask->offset);
  a.fv = make_bitmap_ref(mask->validity, mask->offset);
  a.vv = make_bitmap_ref(values->validity, values->offset);
  a.vbits = make_bitmap_ref(values->data, values->offset);
  a.vdata = bits >= 8 ? reinterpret_cast<const char*>(values->data) + values->offset * (bits / 8) : nullptr;
  a.length = length;
  a.null_sel = null_selection;
  a.super_prefix = pv.super_prefix;
  a.tile_prefix = pv.tile_prefix;
  a.out_data = out_data;
  a.out_valid = reinterpret_cast<uint32_t*>(out_validity);
  a.out_length = &pv.header->out_length;
  a.capacity = out_capacity;
  a.err = ctx->async_err;
  const size_t bitmap_bytes = b200_align_up(size_t((out_capacity + 7) / 8), 4);
  if (out_validity) B200_CUDA(ctx, cudaMemsetAsync(out_validity, 0, bitmap_bytes, ctx->stream));
  if (bits == 1 && out_data) B200_CUDA(ctx, cudaMemsetAsync(out_data, 0, bitmap_bytes, ctx->stream));
  const int64_t nt = n_tiles_for(length);
  switch (bits) {
    case 1: return launch_apply<uint8_t, true>(ctx, a, nt);
    case 8: return launch_apply<uint8_t, false>(ctx, a, nt);
    case 16: return launch_apply<uint16_t, false>(ctx, a, nt);
    case 32: return launch_apply<uint32_t, false>(ctx, a, nt);
    case 64: return launch_apply<uint64_t, false>(ctx, a, nt);
  }
  return b200_fail(ctx, B200_NOT_IMPLEMENTED, "filter: bit width %d", bits);
}

}  // extern "C"
