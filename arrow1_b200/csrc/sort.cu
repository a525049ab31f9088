// SortIndices: stable LSD radix sort of row numbers by an order-preserving 64-bit key.
// Replaces ArraySortIndices::Exec + PartitionNulls + ArrayCompareSorter / ArrayCountSorter
// (reference compute/kernels/vector_sort.cc:224-302, 374-575).
//
//   key(row)  unsigned: v | signed: v ^ sign bit | float: -0.0 -> +0.0, then flip sign bit of
//             non-negatives / all bits of negatives | Descending: bitwise NOT inside the type width.
//             NaN -> 2^64-2 and null -> 2^64-1 in BOTH orders, so the reference's layout
//             [values | NaNs | nulls], the last two in row order (vector_sort.cc:269-302), falls
//             out of the stable sort with no partition pass.  For 64-bit integers every key value
//             is taken, so nulls ride as 2^64-1 and are split from genuine maxima by a 257th bin
//             in the most significant digit pass (decided by the row's validity bit).
//   kernels   sort_hist_kernel   one read of the column: all 8 digit histograms (shared-memory
//                                atomics, flushed once per CTA)
//             sort_scan_kernel   exclusive digit bases; passes whose digit is constant are
//                                marked skipped (generalises the reference's narrow-range
//                                counting sort, vector_sort.cc:484-525); ping-pong parity
//             sort_pass_kernel   onesweep pass: 4096-key tile per CTA, warp-level match ranking,
//                                per-digit decoupled look-back across tiles, keys/payload staged
//                                through shared memory so global stores are digit-contiguous.
//                                The first executed pass reads the raw column (payload = row
//                                number, no iota pass, vector_sort.cc:568); the last one writes
//                                the uint64 indices.
// Algorithmic HBM bytes per row for a 64-bit column with 8 live digits: 8 (hist) + 8+12 (first)
// + 6*24 + 12+8 (last) = 192 B; compulsory 16.125 B/row.
#include <algorithm>

#include "common.cuh"

namespace {

constexpr int kDigits = 8;
constexpr int kThreads = 256;
constexpr int kItems = 16;
constexpr int kTile = kThreads * kItems;  // 4096
constexpr int kWarps = kThreads / 32;
constexpr int kBinStride = 264;           // >= 257, keeps rows 32-byte aligned
constexpr uint32_t kStatusPartial = 1u << 30;
constexpr uint32_t kStatusInclusive = 2u << 30;
constexpr uint32_t kValueMask = (1u << 30) - 1;
constexpr uint32_t kNullFlag = 1u << 31;
constexpr uint64_t kNanKey = ~uint64_t(0) - 1;
constexpr uint64_t kNullKey = ~uint64_t(0);
constexpr int64_t kMaxRows = (int64_t(1) << 30) - 1;

struct PassInfo {
  int32_t skip;
  int32_t first;
  int32_t last;
  int32_t src;  // ping-pong buffer read by this pass (ignored when first)
};

struct SortParams {
  // key source A: a column
  const void* vdata;   // shifted by the element offset
  BitmapRef vv;
  int32_t type;        // b200_type, or 0 when keys_in is used
  int32_t order;
  int32_t nullbin;     // 64-bit integer column with nulls
  int32_t pad;
  // key source B: ready-made keys
  const uint64_t* keys_in;
  int64_t n;
  int64_t n_tiles;
  uint64_t index_base;
  const uint64_t* gather;  // optional: out_idx[pos] = gather[row]
  uint64_t* out_idx;
  uint64_t* out_keys;      // optional
  uint32_t* hist;          // [8][kBinStride] counts, then exclusive bases
  PassInfo* info;          // [8]
  uint32_t* counters;      // [8] tile tickets
  uint32_t* err;
  uint32_t* state;         // [8][n_tiles][kBinStride]
  uint64_t* keys[2];
  uint32_t* pays[2];
};

__device__ __forceinline__ uint64_t f64_key(uint64_t b) {
  if ((b << 1) > 0xFFE0000000000000ull) return kNanKey;      // NaN: exponent all ones, mantissa != 0
  if (b == 0x8000000000000000ull) b = 0;                      // -0.0 == +0.0 under operator<
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ uint64_t f32_key(uint32_t b) {
  if ((b << 1) > 0xFF000000u) return kNanKey;
  if (b == 0x80000000u) b = 0;
  return (b >> 31) ? uint32_t(~b) : (b | 0x80000000u);
}

// order-preserving key of row `row` (valid rows only)
__device__ __forceinline__ uint64_t value_key(const SortParams& p, int64_t row) {
  uint64_t k, wmask;
  switch (p.type) {
    case B200_UINT8: k = reinterpret_cast<const uint8_t*>(p.vdata)[row]; wmask = 0xFFull; break;
    case B200_INT8: k = reinterpret_cast<const uint8_t*>(p.vdata)[row] ^ 0x80u; wmask = 0xFFull; break;
    case B200_UINT16: k = reinterpret_cast<const uint16_t*>(p.vdata)[row]; wmask = 0xFFFFull; break;
    case B200_INT16: k = reinterpret_cast<const uint16_t*>(p.vdata)[row] ^ 0x8000u; wmask = 0xFFFFull; break;
    case B200_UINT32: k = reinterpret_cast<const uint32_t*>(p.vdata)[row]; wmask = 0xFFFFFFFFull; break;
    case B200_INT32: k = reinterpret_cast<const uint32_t*>(p.vdata)[row] ^ 0x80000000u; wmask = 0xFFFFFFFFull; break;
    case B200_FLOAT: {
      k = f32_key(reinterpret_cast<const uint32_t*>(p.vdata)[row]);
      if (k == kNanKey) return k;
      wmask = 0xFFFFFFFFull;
      break;
    }
    case B200_UINT64: k = reinterpret_cast<const uint64_t*>(p.vdata)[row]; wmask = ~0ull; break;
    case B200_INT64: k = reinterpret_cast<const uint64_t*>(p.vdata)[row] ^ 0x8000000000000000ull; wmask = ~0ull; break;
    default: {  // B200_DOUBLE
      k = f64_key(reinterpret_cast<const uint64_t*>(p.vdata)[row]);
      if (k == kNanKey) return k;
      wmask = ~0ull;
      break;
    }
  }
  // Descending: ties keep ascending row order, so invert the key, not the output
  // (vector_sort.cc:394-401, golden vector_sort_test.cc:457-458)
  return p.order ? (~k & wmask) : k;
}

__device__ __forceinline__ uint64_t source_key(const SortParams& p, int64_t row, bool* is_null) {
  if (p.keys_in) {
    *is_null = false;
    return p.keys_in[row];
  }
  const bool valid = bitmap_get(p.vv, row);
  *is_null = !valid;
  return valid ? value_key(p, row) : kNullKey;
}

// Digit histograms without shared-memory atomics (ATOMS costs ~3 cycles per lane on sm_100a):
// every warp owns private histograms; lanes holding the same digit are found with MATCH.ANY and
// the lowest such lane adds the whole group's count with a plain read-modify-write.
struct HistSmem {
  uint32_t h[kWarps][kDigits][kBinStride];
};

__global__ void __launch_bounds__(kThreads) sort_hist_kernel(SortParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  HistSmem& s = *reinterpret_cast<HistSmem*>(smem_raw);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  for (int i = t; i < kWarps * kDigits * kBinStride; i += kThreads) (&s.h[0][0][0])[i] = 0;
  __syncthreads();
  uint32_t(*myh)[kBinStride] = s.h[warp];
  for (int64_t base = int64_t(blockIdx.x) * kTile; base < p.n; base += int64_t(gridDim.x) * kTile) {
#pragma unroll 2
    for (int k = 0; k < kItems; ++k) {
      const int64_t row = base + k * kThreads + t;
      const bool in = row < p.n;
      bool is_null = false;
      const uint64_t key = in ? source_key(p, row, &is_null) : 0;
#pragma unroll
      for (int d = 0; d < kDigits; ++d) {
        uint32_t bin = uint32_t(key >> (8 * d)) & 255u;
        if (d == kDigits - 1 && p.nullbin && is_null) bin = 256;
        if (!in) bin = 0xFFFFu;
        const unsigned m = __match_any_sync(0xffffffffu, bin);
        if (in && lane == __ffs(m) - 1) myh[d][bin] += __popc(m);
      }
      __syncwarp();
    }
  }
  __syncthreads();
  for (int i = t; i < kDigits * kBinStride; i += kThreads) {
    uint32_t c = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) c += (&s.h[w][0][0])[i];
    if (c) atomicAdd(p.hist + i, c);
  }
}

// one warp per digit: exclusive scan of its 257 bins; then thread 0 plans the passes
__global__ void __launch_bounds__(kThreads) sort_scan_kernel(SortParams p) {
  __shared__ int skip_s[kDigits];
  const int lane = threadIdx.x & 31, d = threadIdx.x >> 5;
  uint32_t* h = p.hist + d * kBinStride;
  uint32_t carry = 0;
  bool single = false;
  for (int base = 0; base < 257; base += 32) {
    const int b = base + lane;
    const uint32_t c = b < 257 ? h[b] : 0;
    if (c == uint32_t(p.n)) single = true;
    uint32_t incl = c;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      uint32_t o = __shfl_up_sync(0xffffffffu, incl, s);
      if (lane >= s) incl += o;
    }
    if (b < 257) h[b] = carry + incl - c;
    carry += __shfl_sync(0xffffffffu, incl, 31);
  }
  single = __any_sync(0xffffffffu, single);
  if (lane == 0) skip_s[d] = single ? 1 : 0;
  __syncthreads();
  if (threadIdx.x == 0) {
    int live = 0;
    for (int i = 0; i < kDigits; ++i) live += !skip_s[i];
    if (live == 0) skip_s[0] = 0;  // at least one pass must run to emit the indices
    int first = -1, last = -1;
    for (int i = 0; i < kDigits; ++i)
      if (!skip_s[i]) {
        if (first < 0) first = i;
        last = i;
      }
    int src = 0;
    for (int i = 0; i < kDigits; ++i) {
      PassInfo pi;
      pi.skip = skip_s[i];
      pi.first = (i == first);
      pi.last = (i == last);
      pi.src = src;
      p.info[i] = pi;
      // the first pass writes buffer 0; every later executed pass flips
      if (!pi.skip) src = pi.first ? 0 : (src ^ 1);
    }
  }
}

__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_volatile_u32(uint32_t* p, uint32_t v) {
  asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

struct PassSmem {
  uint64_t keys[kTile];
  uint32_t pay[kTile];
  uint32_t whist[kWarps][kBinStride];
  uint32_t tile_count[kBinStride];
  uint32_t tile_prefix[kBinStride];
  uint32_t bin_delta[kBinStride];
  uint32_t warp_sums[kWarps];
  uint32_t tile;
};

template <int NBINS>
__global__ void __launch_bounds__(kThreads, 3) sort_pass_kernel(SortParams p, int d) {
  const PassInfo info = p.info[d];
  if (info.skip) return;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  PassSmem& s = *reinterpret_cast<PassSmem*>(smem_raw);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const unsigned lt = lanemask_lt();
  const bool top_nullbin = (NBINS == 257) && (d == kDigits - 1);

  if (t == 0) s.tile = atomicAdd(p.counters + d, 1u);
  for (int i = t; i < kWarps * kBinStride; i += kThreads) (&s.whist[0][0])[i] = 0;
  __syncthreads();
  const int64_t tile = s.tile;
  const int64_t tile_row0 = tile * kTile;
  const int tile_items = int(min(int64_t(kTile), p.n - tile_row0));
  const uint64_t* keys_src = p.keys[info.src];
  const uint32_t* pays_src = p.pays[info.src];
  uint32_t* state = p.state + (size_t(d) * p.n_tiles + tile) * kBinStride;

  // ---- load: warp-striped, item i of lane l is local index warp*512 + i*32 + l
  uint64_t key[kItems];
  uint32_t nullmask = 0;  // bit i: item i goes to bin 256
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    const int local = warp * (kItems * 32) + i * 32 + lane;
    key[i] = kNullKey;
    if (local < tile_items) {
      const int64_t row = tile_row0 + local;
      if (info.first) {
        bool is_null;
        key[i] = source_key(p, row, &is_null);
        if (top_nullbin && is_null) nullmask |= 1u << i;
      } else {
        key[i] = keys_src[row];
        if (top_nullbin && (uint32_t(key[i] >> 56) == 255u)) {
          if (!bitmap_get(p.vv, pays_src[row])) nullmask |= 1u << i;
        }
      }
    }
  }

  // ---- rank inside the warp (stable: items visited in index order, lanes in lane order)
  uint32_t rank[kItems];
  uint32_t* myhist = s.whist[warp];
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    const int local = warp * (kItems * 32) + i * 32 + lane;
    uint32_t bin = uint32_t(key[i] >> (8 * d)) & 255u;
    if (top_nullbin && ((nullmask >> i) & 1u)) bin = 256;
    if (local >= tile_items) bin = 0xFFFFu;
    const unsigned m = __match_any_sync(0xffffffffu, bin);
    const int leader = __ffs(m) - 1;
    uint32_t old = 0;
    if (lane == leader && bin != 0xFFFFu) {
      old = myhist[bin];
      myhist[bin] = old + __popc(m);
    }
    old = __shfl_sync(0xffffffffu, old, leader);
    rank[i] = old + __popc(m & lt);
    __syncwarp();
  }
  __syncthreads();

  // ---- per bin: exclusive scan over warps, publish the tile's count
  for (int b = t; b < NBINS; b += kThreads) {
    uint32_t acc = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      const uint32_t c = s.whist[w][b];
      s.whist[w][b] = acc;
      acc += c;
    }
    s.tile_count[b] = acc;
    st_volatile_u32(state + b, (tile == 0 ? kStatusInclusive : kStatusPartial) | acc);
  }
  __syncthreads();
  // ---- exclusive prefix over bins (tile-local layout)
  {
    const uint32_t c = s.tile_count[t];  // bins 0..255
    uint32_t incl = c;
#pragma unroll
    for (int sft = 1; sft < 32; sft <<= 1) {
      uint32_t o = __shfl_up_sync(0xffffffffu, incl, sft);
      if (lane >= sft) incl += o;
    }
    if (lane == 31) s.warp_sums[warp] = incl;
    __syncthreads();
    uint32_t wpre = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w)
      if (w < warp) wpre += s.warp_sums[w];
    s.tile_prefix[t] = wpre + incl - c;
    if (NBINS == 257 && t == kThreads - 1) s.tile_prefix[256] = wpre + incl;
  }
  __syncthreads();
  // ---- decoupled look-back, one thread per bin
  const uint32_t* base_hist = p.hist + d * kBinStride;
  for (int b = t; b < NBINS; b += kThreads) {
    uint32_t excl = 0;
    if (tile > 0) {
      const uint32_t* st = state + b - kBinStride;  // predecessor tile
      int64_t back = tile - 1;
      uint32_t spins = 0;
      while (true) {
        const uint32_t v = ld_volatile_u32(st);
        const uint32_t status = v & ~kValueMask;
        if (status == 0) {
          if (++spins > (1u << 26)) {  // never hang the GPU: flag and bail out
            atomicExch(p.err, 1u);
            break;
          }
          continue;
        }
        excl += v & kValueMask;
        if (status == kStatusInclusive || back == 0) break;
        --back;
        st -= kBinStride;
      }
      st_volatile_u32(state + b, kStatusInclusive | ((excl + s.tile_count[b]) & kValueMask));
    }
    s.bin_delta[b] = base_hist[b] + excl - s.tile_prefix[b];
  }
  __syncthreads();

  // ---- scatter into shared memory in tile-sorted order
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    const int local = warp * (kItems * 32) + i * 32 + lane;
    if (local < tile_items) {
      uint32_t bin = uint32_t(key[i] >> (8 * d)) & 255u;
      const bool isnull = top_nullbin && ((nullmask >> i) & 1u);
      if (isnull) bin = 256;
      const uint32_t pos = s.tile_prefix[bin] + myhist[bin] + rank[i];
      const int64_t row = tile_row0 + local;
      uint32_t pay = info.first ? uint32_t(row) : pays_src[row];
      if (isnull) pay |= kNullFlag;
      s.keys[pos] = key[i];
      s.pay[pos] = pay;
    }
  }
  __syncthreads();

  // ---- digit-contiguous global stores
  uint64_t* keys_dst = p.keys[info.first ? 0 : (info.src ^ 1)];
  uint32_t* pays_dst = p.pays[info.first ? 0 : (info.src ^ 1)];
#pragma unroll 4
  for (int k = 0; k < kItems; ++k) {
    const int j = k * kThreads + t;
    if (j < tile_items) {
      const uint64_t kk = s.keys[j];
      uint32_t pay = s.pay[j];
      uint32_t bin = uint32_t(kk >> (8 * d)) & 255u;
      if (top_nullbin && (pay & kNullFlag)) bin = 256;
      pay &= ~kNullFlag;
      const uint32_t pos = uint32_t(j) + s.bin_delta[bin];
      if (info.last) {
        p.out_idx[pos] = p.gather ? p.gather[pay] : (uint64_t(pay) + p.index_base);
        if (p.out_keys) p.out_keys[pos] = kk;
      } else {
        keys_dst[pos] = kk;
        pays_dst[pos] = pay;
      }
    }
  }
}

__global__ void sort_keys_kernel(SortParams p, uint64_t* out) {
  const int64_t row = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (row < p.n) {
    bool is_null;
    out[row] = source_key(p, row, &is_null);
  }
}

struct Layout {
  size_t hist, info, counters, err, state, zero_bytes, keys0, keys1, pays0, pays1, total;
};

Layout layout_for(int64_t n) {
  Layout L;
  const size_t n_tiles = size_t((n + kTile - 1) / kTile);
  size_t off = 0;
  L.hist = off; off += b200_align_up(size_t(kDigits) * kBinStride * 4, 256);
  L.info = off; off += 256;
  L.counters = off; off += 128;
  L.err = off; off += 128;
  L.state = off; off += b200_align_up(size_t(kDigits) * n_tiles * kBinStride * 4, 256);
  L.zero_bytes = off;
  L.keys0 = off; off += b200_align_up(size_t(n) * 8, 256);
  L.keys1 = off; off += b200_align_up(size_t(n) * 8, 256);
  L.pays0 = off; off += b200_align_up(size_t(n) * 4, 256);
  L.pays1 = off; off += b200_align_up(size_t(n) * 4, 256);
  L.total = off + 256;
  return L;
}

int fill_column_source(b200_ctx* ctx, const b200_col* values, int order, SortParams* p) {
  if (!values) return b200_fail(ctx, B200_INVALID, "sort_indices: null column");
  const int bits = b200_type_bits(values->type);
  // array_sort_indices has kernels for numeric types only (vector_sort.cc:577-606)
  if (bits < 8) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "sort_indices: no kernel for type %d", values->type);
  if (order != 0 && order != 1) return b200_fail(ctx, B200_INVALID, "sort_indices: bad order %d", order);
  if (values->length < 0 || values->offset < 0) return b200_fail(ctx, B200_INVALID, "sort_indices: negative length/offset");
  if (values->length > 0 && !values->data) return b200_fail(ctx, B200_INVALID, "sort_indices: null data buffer");
  const bool has_nulls = values->validity != nullptr && values->null_count != 0;
  p->vdata = reinterpret_cast<const char*>(values->data) + values->offset * (bits / 8);
  p->vv = has_nulls ? make_bitmap_ref(values->validity, values->offset) : BitmapRef{nullptr, 0};
  p->type = values->type;
  p->order = order;
  p->nullbin = (has_nulls && (values->type == B200_INT64 || values->type == B200_UINT64)) ? 1 : 0;
  p->keys_in = nullptr;
  p->n = values->length;
  return B200_OK;
}

int run_sort(b200_ctx* ctx, SortParams p, void* scratch, size_t scratch_bytes) {
  if (p.n == 0) return B200_OK;
  if (p.n > kMaxRows)
    return b200_fail(ctx, B200_NOT_IMPLEMENTED, "sort: %lld rows exceed the per-call limit of 2^30-1; shard the column",
                     (long long)p.n);
  const Layout L = layout_for(p.n);
  if (!scratch || scratch_bytes < L.total) return b200_fail(ctx, B200_INVALID, "sort: scratch too small (%zu < %zu)", scratch_bytes, L.total);
  char* base = reinterpret_cast<char*>(b200_align_up(reinterpret_cast<uintptr_t>(scratch), 256));
  p.n_tiles = (p.n + kTile - 1) / kTile;
  p.hist = reinterpret_cast<uint32_t*>(base + L.hist);
  p.info = reinterpret_cast<PassInfo*>(base + L.info);
  p.counters = reinterpret_cast<uint32_t*>(base + L.counters);
  p.err = reinterpret_cast<uint32_t*>(base + L.err);
  p.state = reinterpret_cast<uint32_t*>(base + L.state);
  p.keys[0] = reinterpret_cast<uint64_t*>(base + L.keys0);
  p.keys[1] = reinterpret_cast<uint64_t*>(base + L.keys1);
  p.pays[0] = reinterpret_cast<uint32_t*>(base + L.pays0);
  p.pays[1] = reinterpret_cast<uint32_t*>(base + L.pays1);
  B200_CUDA(ctx, cudaMemsetAsync(base, 0, L.zero_bytes, ctx->stream));
  const int hist_blocks = int(std::min<int64_t>(p.n_tiles, int64_t(ctx->num_sms) * 3));
  static bool attr_set = false;
  const int smem = int(sizeof(PassSmem));
  const int hist_smem = int(sizeof(HistSmem));
  if (!attr_set) {
    B200_CUDA(ctx, cudaFuncSetAttribute(sort_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, hist_smem));
    B200_CUDA(ctx, cudaFuncSetAttribute(sort_pass_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    B200_CUDA(ctx, cudaFuncSetAttribute(sort_pass_kernel<257>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  sort_hist_kernel<<<hist_blocks, kThreads, hist_smem, ctx->stream>>>(p);
  B200_LAUNCHED(ctx);
  sort_scan_kernel<<<1, kThreads, 0, ctx->stream>>>(p);
  B200_LAUNCHED(ctx);
  for (int d = 0; d < kDigits; ++d) {
    if (p.nullbin && d == kDigits - 1)
      sort_pass_kernel<257><<<(unsigned)p.n_tiles, kThreads, smem, ctx->stream>>>(p, d);
    else
      sort_pass_kernel<256><<<(unsigned)p.n_tiles, kThreads, smem, ctx->stream>>>(p, d);
    B200_LAUNCHED(ctx);
  }
  return B200_OK;
}

}  // namespace

extern "C" {

int b200_sort_indices_scratch_bytes(int64_t length, size_t* out_bytes) {
  if (!out_bytes || length < 0) return B200_INVALID;
  *out_bytes = layout_for(length).total + 256;
  return B200_OK;
}

int b200_sort_pairs_scratch_bytes(int64_t length, size_t* out_bytes) {
  return b200_sort_indices_scratch_bytes(length, out_bytes);
}

int b200_sort_indices(b200_ctx* ctx, const b200_col* values, int order, uint64_t index_base,
                      uint64_t* out_indices, void* scratch, size_t scratch_bytes) {
  if (!ctx) return B200_INVALID;
  SortParams p;
  memset(&p, 0, sizeof(p));
  B200_RETURN_NOT_OK(fill_column_source(ctx, values, order, &p));
  if (p.n > 0 && !out_indices) return b200_fail(ctx, B200_INVALID, "sort_indices: null output");
  p.index_base = index_base;
  p.out_idx = out_indices;
  return run_sort(ctx, p, scratch, scratch_bytes);
}

int b200_sort_keys(b200_ctx* ctx, const b200_col* values, int order, uint64_t* out_keys) {
  if (!ctx) return B200_INVALID;
  SortParams p;
  memset(&p, 0, sizeof(p));
  B200_RETURN_NOT_OK(fill_column_source(ctx, values, order, &p));
  if (p.n == 0) return B200_OK;
  if (!out_keys) return b200_fail(ctx, B200_INVALID, "sort_keys: null output");
  sort_keys_kernel<<<(unsigned)((p.n + 255) / 256), 256, 0, ctx->stream>>>(p, out_keys);
  B200_LAUNCHED(ctx);
  return B200_OK;
}

int b200_sort_pairs(b200_ctx* ctx, const uint64_t* keys, const uint64_t* payload, int64_t length,
                    uint64_t* out_keys, uint64_t* out_payload, void* scratch, size_t scratch_bytes) {
  if (!ctx) return B200_INVALID;
  if (length < 0) return b200_fail(ctx, B200_INVALID, "sort_pairs: negative length");
  if (length == 0) return B200_OK;
  if (!keys || !out_payload) return b200_fail(ctx, B200_INVALID, "sort_pairs: null argument");
  SortParams p;
  memset(&p, 0, sizeof(p));
  p.keys_in = keys;
  p.n = length;
  p.gather = payload;  // NULL => emits the permutation itself
  p.out_idx = out_payload;
  p.out_keys = out_keys;
  return run_sort(ctx, p, scratch, scratch_bytes);
}

}  // extern "C"
