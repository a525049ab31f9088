// SortIndices: stable LSD radix sort of row numbers by an order-preserving unsigned key.
// Replaces ArraySortIndices::Exec + PartitionNulls + ArrayCompareSorter / ArrayCountSorter
// (reference compute/kernels/vector_sort.cc:224-302, 374-575).
//
//   key(row)  unsigned: v | signed: v ^ sign bit | float: -0.0 -> +0.0, then flip the sign bit of
//             non-negatives / all bits of negatives | Descending: bitwise NOT inside the type width
//             (ties keep ascending row order in both orders, vector_sort.cc:394-401).
//   classes   the reference's layout [values | NaNs | nulls], the last two in row order for both
//             orders (vector_sort.cc:269-302), costs no partition pass and no special bins: NaN and null rows
//             ride through every pass with an all-ones key, so the stable sort leaves them — together with
//             genuine maxima — at the very end in row order, and tail_*_kernel split that tail (its length is
//             counted by the first pass) stably into [maxima | NaNs | nulls]; null-ness is read from the
//             validity bit, so a genuine INT64_MAX is never confused with a null.
//   kernels   sort_hist_kernel   one read of the column: the digit histograms of every live digit
//                                (shared-memory atomics, flushed once per CTA)
//             sort_scan_kernel   exclusive digit bases; a digit that is constant over the column is
//                                skipped (generalises the reference's narrow-range counting sort,
//                                vector_sort.cc:484-525; also makes int32 a 4-pass sort)
//             sort_pass_kernel   onesweep pass: 4096-key tile per CTA; lanes with equal digits are
//                                matched with one warp ballot per digit bit (MATCH.ANY measured ~60
//                                cycles/warp-instruction on B200; an atomicOr bitmask in shared memory
//                                cost three dependent shared round trips per item); per-digit decoupled
//                                look-back across tiles; keys + 32-bit row numbers staged through
//                                shared memory so global stores are digit-contiguous.  The first
//                                live pass reads the raw column (row number = position, no iota
//                                pass, vector_sort.cc:568); the last one writes the uint64 indices.
// Algorithmic HBM bytes per row, 64-bit column with 8 live digits: 8.125 (hist) + 8.125+12 (first)
// + 6*24 + 12+8 (last) = 192 B; compulsory 16.125 B/row.
#include <stdlib.h>

#include <algorithm>
#include <type_traits>
#include <vector>
#include "common.cuh"

namespace {

constexpr int kDigits = 8;
constexpr int kHistThreads = 256;          // histogram / scan kernels (the scan wants one warp per digit)
constexpr int kHistItems = 16;
constexpr int kHistTile = kHistThreads * kHistItems;
#ifndef B200_SORT_THREADS
#define B200_SORT_THREADS 256
#define B200_SORT_ITEMS 16
#define B200_SORT_MINBLOCKS 3
#endif
#ifndef B200_SORT_NARROW_MINBLOCKS
#define B200_SORT_NARROW_MINBLOCKS (B200_SORT_MINBLOCKS + 1)   // 32-bit keys: 64 registers, 4 CTAs per SM
#endif
constexpr int kThreads = B200_SORT_THREADS;   // pass kernel: threads x keys per thread (tuned on B200, see
constexpr int kItems = B200_SORT_ITEMS;       // profiles/r01_sort_tile_sweep.txt)
constexpr int kTile = kThreads * kItems;
constexpr int kWarps = kThreads / 32;
constexpr int kMinBlocks = B200_SORT_MINBLOCKS;
static_assert(kThreads >= 256 && kItems <= 32, "one thread per digit bin; validity words come from lanes < kItems");
constexpr int kBinStride = 264;           // >= 258, keeps rows 32-byte aligned
constexpr int kLook = 8;                  // predecessor states requested per look-back round trip
constexpr uint32_t kStatusPartial = 1u << 30;
constexpr uint32_t kStatusInclusive = 2u << 30;
constexpr uint32_t kValueMask = (1u << 30) - 1;
constexpr uint64_t kNanKeyOut = ~uint64_t(0) - 1;   // b200_sort_keys: NaN
constexpr uint64_t kNullKeyOut = ~uint64_t(0);      // b200_sort_keys: null
constexpr int64_t kMaxRows = (int64_t(1) << 30) - 1;
enum { kUnsignedKey = 0, kSignedKey = 1, kFloatKey = 2 };

struct PassInfo {
  int32_t skip;
  int32_t first;
  int32_t last;
  int32_t src;         // ping-pong buffer read by this pass (ignored when first)
  int32_t out_narrow;  // the pass writes 32-bit keys: only the word the remaining passes still look at
  int32_t hw;          // high-word-first plan (see HwState): the last pass also leaves the sorted high words behind
  int32_t pad[2];
};

// ---- high word first (64-bit keys whose upper 32 bits almost identify the row) ---------------------------------------
// An LSD sort of 64-bit keys runs 8 passes although, for n << 2^32 keys with a high-entropy upper word (ids, hashes,
// random int64), the upper word alone orders all but a few percent of the rows.  The plan kernel estimates the number of
// distinct upper words from the digit histograms it already has (2^(sum of the entropies of digits 4..7), an upper bound
// of the joint figure) and, when that is >= 8 n, runs ONLY the four narrow passes over digits 4..7 — stable, so rows
// that tie on the upper word stay in row order — and then repairs the ties: every run of equal upper words (found in
// the sorted upper words the last pass leaves behind) is sorted by the lower word, read from the column, by one
// thread (runs of up to 32 rows).  The run of the all-ones upper word holds the null / NaN rows (key all ones): it is
// first split stably into [other keys | genuine maxima | NaNs | nulls] by the tail kernels, and only its first part is a
// run to repair.  If a run is longer than 32 rows the estimate was wrong (correlated digits): a flag is raised and the
// ordinary 8-pass plan runs after all (round 2: its launches are gated on the flag and otherwise exit at once).
constexpr int kHwMaxRun = 32;   // longest run of equal upper words the repair sorts (one thread per run)
struct HwState {
  uint32_t hw;            // the plan kernel chose high word first
  uint32_t fail;          // a tie run was too long: round 2 runs the ordinary plan
  int64_t top_start;      // first position of the run of the all-ones upper word (n if there is none)
  uint32_t class_total[4];
  uint32_t n_runs;        // runs of equal upper words found by the repair
  uint32_t top_count;     // keys whose upper word is all ones (counted by the histogram kernel)
};

struct SortParams {
  // key source A: a column
  const void* vdata;   // shifted by the element offset
  BitmapRef vv;        // validity (words == nullptr: no nulls)
  int32_t width;       // bytes per value
  int32_t tkind;       // kUnsignedKey / kSignedKey / kFloatKey
  int32_t cls_digit;   // >= 0: the column can hold nulls or NaNs (its most significant digit), -1 = it cannot
  int32_t n_digits;    // digits that can vary (width, or 8 for ready-made keys)
  uint64_t sign;       // sign bit of the type
  uint64_t wmask;      // all ones inside the type width
  uint64_t inv;        // wmask for Descending, else 0
  uint64_t nan_above;  // float: (bits & ~sign) > nan_above  <=>  NaN
  // key source B: ready-made keys
  const uint64_t* keys_in;
  int64_t n;
  int64_t n_tiles;
  uint64_t index_base;
  const uint64_t* gather;  // optional: out_idx[pos] = gather[row]
  const uint32_t* pay_in;  // optional (with keys_in): 32-bit payload carried through the passes instead of the row number
  uint64_t* out_idx;
  uint64_t* out_keys;      // optional
  int32_t tail_fix;        // NaN / null rows are ordered by the tail fix-up after the last pass (no 258-bin class pass)
  uint32_t* tail_count;    // rows whose key is all ones: nulls, NaNs and genuine maxima (they end up last, in row order)
  int32_t narrow;          // passes that only look at one 32-bit word of the key carry that word alone
  uint32_t* hist;          // [8][kBinStride] counts, then exclusive bases
  PassInfo* info;          // [8]
  PassInfo* info2;         // [8] the ordinary plan, for round 2 of a failed high-word-first attempt
  HwState* hws;
  int32_t hw_allowed;      // host: a 64-bit column, indices only
  int32_t round2;          // this launch belongs to round 2: it runs only if hws->fail is set
  uint32_t* counters;      // [8] tile tickets
  uint32_t* err;
  uint32_t* state;         // [8][n_tiles][kBinStride]
  uint64_t* keys[2];
  uint32_t* pays[2];
};

// raw (zero-extended) value bits -> key; *cls = 1 for NaN
__device__ __forceinline__ uint64_t transform_key(const SortParams& p, uint64_t raw, uint32_t* cls) {
  uint64_t k = raw;
  *cls = 0;
  if (p.tkind == kSignedKey) {
    k ^= p.sign;
  } else if (p.tkind == kFloatKey) {
    if ((k & ~p.sign) > p.nan_above) {
      *cls = 1;
      return p.wmask;
    }
    if (k == p.sign) k = 0;  // -0.0 == +0.0 under operator<
    k = (k & p.sign) ? (~k & p.wmask) : (k | p.sign);
  }
  return k ^ p.inv;
}

__device__ __forceinline__ uint64_t load_raw(const SortParams& p, int64_t row) {
  switch (p.width) {
    case 1: return reinterpret_cast<const uint8_t*>(p.vdata)[row];
    case 2: return reinterpret_cast<const uint16_t*>(p.vdata)[row];
    case 4: return reinterpret_cast<const uint32_t*>(p.vdata)[row];
    default: return reinterpret_cast<const uint64_t*>(p.vdata)[row];
  }
}

// key and class (0 value, 1 NaN, 2 null) of a source row
__device__ __forceinline__ uint64_t source_key(const SortParams& p, int64_t row, uint32_t* cls) {
  if (p.keys_in) {
    *cls = 0;
    return p.keys_in[row];
  }
  if (!bitmap_get(p.vv, row)) {
    *cls = 2;
    return p.wmask;
  }
  return transform_key(p, load_raw(p, row), cls);
}

// All digit histograms in one read of the column: shared-memory atomics, flushed once per CTA.
// (Measured on B200 at 100 M 64-bit keys: 0.4 ms this way vs 5.3 ms with MATCH.ANY-aggregated
// warp-private counters.)
// kWide: 64-bit keys (a 64-bit column or ready-made keys) with all 8 digits live; its own instantiation, so that the
// generic rows of the other shapes keep their 26 registers and 8 CTAs per SM
template <bool kWide>
__global__ void __launch_bounds__(kHistThreads) sort_hist_kernel(SortParams p) {
  __shared__ uint32_t h[kDigits][kBinStride];
  __shared__ uint32_t top_s;   // keys whose upper word is all ones
  const int t = threadIdx.x;
  for (int i = t; i < kDigits * kBinStride; i += kHistThreads) (&h[0][0])[i] = 0;
  if (t == 0) top_s = 0;
  __syncthreads();
  const int lane = t & 31, warp = t >> 5;
  uint32_t top_mine = 0;
  const bool wide = kWide;
  for (int64_t base = int64_t(blockIdx.x) * kHistTile; base < p.n; base += int64_t(gridDim.x) * kHistTile) {
    if (wide && base + kHistTile <= p.n) {
      // full tile of 64-bit keys (the generic rows below cost ~120 instructions each: type switch, bitmap probe and
      // 64-bit shifts per digit): a warp takes 512 consecutive rows, all 16 loads of a lane issued together, one
      // validity word per 32 rows, digits cut out of the two key halves with PRMT
      const int64_t wrow0 = base + int64_t(warp) * (kHistItems * 32);
      uint32_t vword = 0xffffffffu;   // validity of rows wrow0 + 32 * lane .. + 31 (lanes 0..15)
      if (p.vv.words && !p.keys_in && lane < kHistItems) vword = bitmap_load32(p.vv, wrow0 + lane * 32, p.n);
      uint64_t raw[kHistItems];
      const uint64_t* src = p.keys_in ? p.keys_in : reinterpret_cast<const uint64_t*>(p.vdata);
#pragma unroll
      for (int k = 0; k < kHistItems; ++k) raw[k] = src[wrow0 + k * 32 + lane];
#pragma unroll
      for (int k = 0; k < kHistItems; ++k) {
        uint64_t key = raw[k];
        if (!p.keys_in) {
          uint32_t cls;
          key = transform_key(p, raw[k], &cls);
          if (!((__shfl_sync(0xffffffffu, vword, k) >> lane) & 1u)) key = p.wmask;
        }
        const uint32_t lo = uint32_t(key), hi = uint32_t(key >> 32);
        top_mine += hi == 0xFFFFFFFFu ? 1u : 0u;
#pragma unroll
        for (int d = 0; d < 4; ++d) {
          atomicAdd(&h[d][__byte_perm(lo, 0u, 0x4440u | uint32_t(d))], 1u);
          atomicAdd(&h[d + 4][__byte_perm(hi, 0u, 0x4440u | uint32_t(d))], 1u);
        }
      }
      continue;
    }
#pragma unroll 4
    for (int k = 0; k < kHistItems; ++k) {
      const int64_t row = base + k * kHistThreads + t;
      if (row < p.n) {
        uint32_t cls;
        const uint64_t key = source_key(p, row, &cls);
        top_mine += (key >> 32) == 0xFFFFFFFFull ? 1u : 0u;
#pragma unroll
        for (int d = 0; d < kDigits; ++d) {
          if (d < p.n_digits) {
            atomicAdd(&h[d][uint32_t(key >> (8 * d)) & 255u], 1u);
          }
        }
      }
    }
  }
  top_mine = __reduce_add_sync(0xffffffffu, top_mine);
  if (lane == 0 && top_mine) atomicAdd(&top_s, top_mine);
  __syncthreads();
  for (int i = t; i < kDigits * kBinStride; i += kHistThreads) {
    const uint32_t c = (&h[0][0])[i];
    if (c) atomicAdd(p.hist + i, c);
  }
  if (t == 0 && top_s && p.hws) atomicAdd(&p.hws->top_count, top_s);
}

// the passes of a plan: which digits run, which buffers they read, where the keys become 32-bit
__device__ void write_plan(const SortParams& p, const int* skip_in, bool hw, PassInfo* out) {
  int skip[kDigits];
  int live = 0;
  for (int i = 0; i < kDigits; ++i) {
    skip[i] = skip_in[i] || (hw && i < 4);
    live += !skip[i];
  }
  if (live == 0) skip[0] = 0;  // at least one pass must run to emit the indices
  int first = -1, last = -1;
  for (int i = 0; i < kDigits; ++i)
    if (!skip[i]) {
      if (first < 0) first = i;
      last = i;
    }
  int src = 0;
  for (int i = 0; i < kDigits; ++i) {
    PassInfo pi;
    pi.skip = skip[i];
    pi.first = (i == first);
    pi.last = (i == last);
    pi.src = src;
    // 32-bit keys from the point where every pass still to run looks at the same word: the upper word of a
    // 64-bit key after the last live pass over digits 0..3, the only word of a <= 32-bit column from the start
    int next = -1;
    for (int j = i + 1; j < kDigits; ++j)
      if (!skip[j]) {
        next = j;
        break;
      }
    pi.out_narrow = p.narrow && next >= 0 && (next >= 4 || p.n_digits <= 4);
    pi.hw = hw ? 1 : 0;
    pi.pad[0] = pi.pad[1] = 0;
    out[i] = pi;
    // the first pass writes buffer 0; every later executed pass flips
    if (!pi.skip) src = pi.first ? 0 : (src ^ 1);
  }
}

// one warp per digit: exclusive scan of its 258 bins (and the digit's entropy); then thread 0 plans the passes
__global__ void __launch_bounds__(kHistThreads) sort_scan_kernel(SortParams p) {
  __shared__ int skip_s[kDigits];
  __shared__ double entropy_s[kDigits];
  const int lane = threadIdx.x & 31, d = threadIdx.x >> 5;
  uint32_t* h = p.hist + d * kBinStride;
  uint32_t carry = 0;
  bool single = d >= p.n_digits;  // digits above the type width are constant (zero)
  double clogc = 0.0;             // sum of c * log2(c) over the digit's bins
  for (int base = 0; base < 258; base += 32) {
    const int b = base + lane;
    const uint32_t c = b < 258 ? h[b] : 0;
    if (c == uint32_t(p.n)) single = true;
    if (c > 1) clogc += double(c) * log2(double(c));
    uint32_t incl = c;
#pragma unroll
    for (int s = 1; s < 32; s <<= 1) {
      uint32_t o = __shfl_up_sync(0xffffffffu, incl, s);
      if (lane >= s) incl += o;
    }
    if (b < 258) h[b] = carry + incl - c;
    carry += __shfl_sync(0xffffffffu, incl, 31);
  }
  single = __any_sync(0xffffffffu, single);
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) clogc += __shfl_xor_sync(0xffffffffu, clogc, s);
  if (lane == 0) {
    skip_s[d] = single ? 1 : 0;
    // H = log2(n) - sum(c log2 c) / n bits
    entropy_s[d] = (d < p.n_digits && p.n > 0) ? log2(double(p.n)) - clogc / double(p.n) : 0.0;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    // high word first?  2^(entropies of digits 4..7) bounds the number of distinct upper words from above; with D
    // distinct, equally likely words a row ties with another with probability ~ n / D
    bool hw = false;
    if (p.hw_allowed && p.n_digits == 8 && p.narrow && p.n > 1) {
      const double bits = entropy_s[4] + entropy_s[5] + entropy_s[6] + entropy_s[7];
      hw = bits >= log2(double(p.n)) + 3.0;
      if (p.hw_allowed == 2) hw = !(skip_s[4] && skip_s[5] && skip_s[6] && skip_s[7]);   // forced (tests)
      // the run of the all-ones upper word: a column sort splits it with the tail kernels (nulls, NaNs); plain
      // (key, payload) pairs have no such help, and a long run there would only make the repair give up
      if (!p.tail_fix && p.hws->top_count > uint32_t(kHwMaxRun)) hw = false;
    }
    write_plan(p, skip_s, hw, p.info);
    if (p.hws) {
      p.hws->hw = hw ? 1u : 0u;
      p.hws->top_start = p.n;
      if (p.info2) write_plan(p, skip_s, false, p.info2);
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

// byte `d` of a 64-bit key: word select + one PRMT (the runtime 64-bit shift costs a funnel shift and a mask)
__device__ __forceinline__ uint32_t key_digit(uint64_t key, int d) {
  const uint32_t w = d >= 4 ? uint32_t(key >> 32) : uint32_t(key);
  return __byte_perm(w, 0u, 0x4440u | uint32_t(d & 3));
}

__device__ __forceinline__ uint32_t key_digit(uint32_t key, int d) { return __byte_perm(key, 0u, 0x4440u | uint32_t(d & 3)); }

// m &= lanes whose bit `b` of `bin` equals this lane's: predicate, ballot, conditional complement, and
__device__ __forceinline__ unsigned match_bit(unsigned m, uint32_t bin, uint32_t bitmask) {
  unsigned r;
  asm volatile(
      "{\n"
      "  .reg .pred p;\n"
      "  .reg .b32 t;\n"
      "  and.b32 t, %2, %3;\n"
      "  setp.ne.u32 p, t, 0;\n"
      "  vote.sync.ballot.b32 t, p, 0xffffffff;\n"
      "  @!p not.b32 t, t;\n"
      "  and.b32 %0, %1, t;\n"
      "}\n"
      : "=r"(r)
      : "r"(m), "r"(bin), "r"(bitmask));
  return r;
}

// ---- bulk-asynchronous tile loads (sm_90+ cp.async.bulk, the 1-D form of TMA; SASS: UBLKCP + SYNCS) -----------------
// A pass after the first reads its tile as two contiguous ranges (4096 keys, 4096 row numbers).  With B200_SORT_TMA one
// thread hands both to the copy engine as soon as it knows the tile — before the CTA has even zeroed its histograms —
// and the engine lands them in the staging arrays the scatter will later overwrite; the threads then read their 16
// keys and row numbers with conflict-free shared loads instead of 32 global loads each.
// MEASURED AND LEFT OFF (round 2, profiles/r02_sort_variants.txt): 5.89 ms per 100 M int64 rows against 5.60 ms with
// the per-thread LDGs.  A CTA handles one tile, so nothing is prefetched: the bulk copy only changes who asks for the
// bytes, makes all 256 threads wait for the whole 48 KB instead of each for its own 32 requests, and adds the shared
// loads.  The build keeps it behind the macro (-DB200_SORT_TMA=1) for the persistent, double-buffered form.
#ifndef B200_SORT_TMA
#define B200_SORT_TMA 0
#endif
// lanes of a warp that hold the same digit: matched with one vote per digit bit (0), or through one shared-memory word
// per (warp, digit) — ATOMS.OR, LDS, STS per item instead of 8 VOTE + 24 ALU instructions (1).  MEASURED AND LEFT OFF
// (round 2, profiles/r02_sort_variants.txt): 5.85 ms per 100 M int64 rows against 5.66 ms with the votes — the pass
// already spends its shared-memory wavefronts on digit-random addresses (48 % bank conflicts), three more per item
// cost more than the issue slots they free.
// digits 0..3 launched in the looping form when the high-word-first plan may skip them (see run_sort).  MEASURED AND
// LEFT OFF (profiles/r02_sort_variants.txt): a skipped pass costs 4 us instead of 17 (sort_i64 3.15 -> 3.10 ms), but a
// column that keeps the 8-pass plan runs those four passes in the looping form, which spills more: sort_f64 5.52 -> 6.04 ms
#ifndef B200_SORT_LOOP_LOW
#define B200_SORT_LOOP_LOW 0
#endif
#ifndef B200_SORT_MATCH_ATOMIC
#define B200_SORT_MATCH_ATOMIC 0
#endif
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return uint32_t(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// false if the bytes never arrived (never hang the GPU: the caller flags the context)
__device__ __forceinline__ bool mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  for (uint32_t spins = 0; spins < (1u << 24); ++spins) {
    uint32_t ok;
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok)
                 : "r"(addr), "r"(parity)
                 : "memory");
    if (ok) return true;
  }
  return false;
}

template <typename KeyT>
struct PassSmem {
  KeyT keys[kTile];
  uint32_t pay[kTile];
  uint32_t whist[kWarps][kBinStride];
  uint32_t tile_count[kBinStride];
  uint32_t tile_prefix[kBinStride];
  uint32_t bin_delta[kBinStride];
  uint32_t warp_sums[kWarps];
  uint32_t tile;
  uint64_t mbar;   // completion barrier of the tile's bulk loads
};

// kFull: the tile has 4096 rows.  (NaN / null rows are no business of the passes: they carry the all-ones key and
// are ordered by the tail fix-up after the last pass; round 1 ranked every row of the last pass over 258 bins.)
// kNarrow: the pass looks at one 32-bit word of the key only (digits 4..7 of a 64-bit key, or any digit of a
// <= 32-bit column) and holds, stages and writes that word alone: 8 B per row and pass instead of 12.
// kPairs: (key, payload) pairs under a possible high-word-first plan (its own instantiation: the column sorts' kernels
// stay as they were)
template <bool kFull, bool kNarrow, bool kPairs>
__device__ __forceinline__ void sort_pass_body(const SortParams& p, const int d, const PassInfo info,
                                               PassSmem<typename std::conditional<kNarrow, uint32_t, uint64_t>::type>& s,
                                               const int64_t tile, const int tile_items) {
  using KeyT = typename std::conditional<kNarrow, uint32_t, uint64_t>::type;
  constexpr int NBINS = 256;
  const int word_shift = (kNarrow && d >= 4) ? 32 : 0;                    // which word a narrow key is
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const unsigned lt = lanemask_lt();
  const int64_t tile_row0 = tile * kTile;
  const KeyT* keys_src = reinterpret_cast<const KeyT*>(p.keys[info.src]);
  const uint32_t* pays_src = p.pays[info.src];
  uint32_t* state = p.state + (size_t(d) * p.n_tiles + tile) * kBinStride;
  const int wbase = warp * (kItems * 32) + lane;  // item i sits at local index wbase + 32*i

  // ---- load: all global loads are issued back to back, transforms follow
  KeyT key[kItems];
  if (info.first && !p.keys_in) {
    uint32_t vword = 0xffffffffu;  // validity word of step `lane` (lanes 0..15)
    if (p.vv.words && lane < kItems) vword = bitmap_load32(p.vv, tile_row0 + warp * (kItems * 32) + lane * 32, p.n);
    const char* src = reinterpret_cast<const char*>(p.vdata);
    uint64_t raw[kItems];
    uint32_t tail = 0;
    switch (p.width) {
      case 1:
#pragma unroll
        for (int i = 0; i < kItems; ++i)
          raw[i] = (kFull || wbase + 32 * i < tile_items) ? reinterpret_cast<const uint8_t*>(src)[tile_row0 + wbase + 32 * i] : 0;
        break;
      case 2:
#pragma unroll
        for (int i = 0; i < kItems; ++i)
          raw[i] = (kFull || wbase + 32 * i < tile_items) ? reinterpret_cast<const uint16_t*>(src)[tile_row0 + wbase + 32 * i] : 0;
        break;
      case 4:
#pragma unroll
        for (int i = 0; i < kItems; ++i)
          raw[i] = (kFull || wbase + 32 * i < tile_items) ? reinterpret_cast<const uint32_t*>(src)[tile_row0 + wbase + 32 * i] : 0;
        break;
      default:
#pragma unroll
        for (int i = 0; i < kItems; ++i)
          raw[i] = (kFull || wbase + 32 * i < tile_items) ? reinterpret_cast<const uint64_t*>(src)[tile_row0 + wbase + 32 * i] : 0;
        break;
    }
#pragma unroll
    for (int i = 0; i < kItems; ++i) {
      const uint32_t vw = __shfl_sync(0xffffffffu, vword, i);
      uint32_t cls;
      uint64_t k = transform_key(p, raw[i], &cls);
      if (!((vw >> lane) & 1u)) {
        cls = 2;
        k = p.wmask;
      }
      key[i] = KeyT(k >> word_shift);
      if (k == p.wmask && (kFull || wbase + 32 * i < tile_items)) ++tail;
    }
    // rows whose key is all ones (nulls, NaNs, genuine maxima): the tail the fix-up kernels re-order after the last pass
    if (p.tail_fix) {
      tail = __reduce_add_sync(0xffffffffu, tail);
      if (lane == 0 && tail) atomicAdd(p.tail_count, tail);
    }
  } else {
    if (info.first) {   // ready-made 64-bit keys
#pragma unroll
      for (int i = 0; i < kItems; ++i)
        key[i] = (kFull || wbase + 32 * i < tile_items) ? KeyT(p.keys_in[tile_row0 + wbase + 32 * i] >> word_shift) : KeyT(0);
    } else if (kFull && B200_SORT_TMA) {
      // the tile was requested by the kernel's prologue: wait for the copy engine, read from the staging arrays
      if (!mbar_wait(&s.mbar, 0)) atomicExch(p.err, 3u);
#pragma unroll
      for (int i = 0; i < kItems; ++i) key[i] = s.keys[wbase + 32 * i];
    } else {
#pragma unroll
      for (int i = 0; i < kItems; ++i) key[i] = (kFull || wbase + 32 * i < tile_items) ? keys_src[tile_row0 + wbase + 32 * i] : KeyT(0);
    }
  }

  // ---- rank inside the warp (stable: items visited in index order, lanes in lane order).
  // Lanes holding the same digit are matched with one ballot per digit bit (no shared-memory round
  // trip); the lowest matched lane bumps the warp's digit counter with one native ATOMS.ADD and its
  // return value is the group's base rank.  Items go in groups of kRankGroup: all atomics of a group
  // are issued before the first return value is consumed, so their latencies overlap instead of
  // forming a 16-deep dependency chain (the short-scoreboard stall that dominated v2's profile).
  constexpr int NB = 8;
  constexpr int kRankGroup = kItems < 8 ? kItems : 8;
  uint32_t rank2[kItems / 2];  // two 16-bit ranks per register
  uint32_t* myhist = s.whist[warp];
  uint32_t* match_words = s.pay + warp * NBINS;   // the staging array is free until the scatter; zeroed by the prologue
#pragma unroll
  for (int g0 = 0; g0 < kItems; g0 += kRankGroup) {
    unsigned peers[kRankGroup];
    uint32_t base[kRankGroup];
#pragma unroll
    for (int q = 0; q < kRankGroup; ++q) {
      const int i = g0 + q;
      const bool in = kFull || wbase + 32 * i < tile_items;
      const uint32_t bin = key_digit(key[i], d);
      unsigned m;
      if (B200_SORT_MATCH_ATOMIC && !B200_SORT_TMA) {
        // lanes holding the same digit find each other through one word per (warp, digit) in shared memory: every
        // lane ORs its bit in, reads the word back, and clears it for the next item — three shared-memory
        // instructions whose addresses do not depend on each other's results, against eight votes and 24 ALU
        // instructions (a warp's shared-memory instructions execute in program order)
        uint32_t* word = match_words + bin;
        if (in) atomicOr(word, 1u << lane);
        __syncwarp();
        m = in ? *reinterpret_cast<volatile uint32_t*>(word) : 0u;
        __syncwarp();
        if (in) *reinterpret_cast<volatile uint32_t*>(word) = 0u;
      } else {
        m = kFull ? 0xffffffffu : __ballot_sync(0xffffffffu, in);
#pragma unroll
        for (int b = 0; b < NB; ++b) m = match_bit(m, bin, 1u << b);
      }
      peers[q] = m;
      base[q] = 0;
      // same-warp shared atomics issue in program order: item i's bump precedes item i+1's
      if (in && (m & lt) == 0) base[q] = atomicAdd(myhist + bin, uint32_t(__popc(m)));
      __syncwarp();
    }
#pragma unroll
    for (int q = 0; q < kRankGroup; ++q) {
      const int i = g0 + q;
      const unsigned m = peers[q];
      const uint32_t b0 = __shfl_sync(0xffffffffu, base[q], (__ffs(m) - 1) & 31);
      const uint32_t r = b0 + __popc(m & lt);
      if (i & 1) rank2[i >> 1] |= r << 16;
      else rank2[i >> 1] = r;
    }
  }
  // ---- row numbers: issued now, consumed after the look-back
  uint32_t pay[kItems];
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    const int local = wbase + 32 * i;
    if (info.first && (!p.pay_in || (kPairs && info.hw))) pay[i] = uint32_t(tile_row0 + local);   // (high word first: the POSITION travels, see below)
    else if (kFull && B200_SORT_TMA && !info.first) pay[i] = s.pay[local];
    else pay[i] = (kFull || local < tile_items) ? (info.first ? p.pay_in : pays_src)[tile_row0 + local] : 0u;
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
  // ---- exclusive prefix over bins (tile-local layout): threads 0..255 own one bin each
  {
    uint32_t c = 0, incl = 0;
    if (t < 256) {
      c = s.tile_count[t];
      incl = c;
#pragma unroll
      for (int sft = 1; sft < 32; sft <<= 1) {
        uint32_t o = __shfl_up_sync(0xffffffffu, incl, sft);
        if (lane >= sft) incl += o;
      }
      if (lane == 31) s.warp_sums[warp] = incl;
    }
    __syncthreads();
    if (t < 256) {
      uint32_t wpre = 0;
#pragma unroll
      for (int w = 0; w < 8; ++w)
        if (w < warp) wpre += s.warp_sums[w];
      const uint32_t pre = wpre + incl - c;
      s.tile_prefix[t] = pre;
      // fold the bin's tile offset into every warp's offset (conflict-free: one bin per thread), so the
      // scatter below needs one random shared load per key instead of two — the pass is bound by
      // shared-memory wavefronts (57 % of them bank conflicts from digit-random addresses)
#pragma unroll
      for (int w = 0; w < kWarps; ++w) s.whist[w][t] += pre;
    }
  }
  __syncthreads();
  // ---- decoupled look-back, one thread per bin.  The states of the kLook preceding tiles are
  // requested together (one L2 round trip instead of one per predecessor: the serial walk was 25 %
  // of the kernel's stall samples) and before the shared-memory scatter, which hides the first trip.
  const uint32_t* base_hist = p.hist + d * kBinStride;
  uint32_t look[kLook];
  if (tile > 0 && t < 256) {
#pragma unroll
    for (int j = 0; j < kLook; ++j)
      look[j] = tile - 1 - j >= 0 ? ld_volatile_u32(state + t - (j + 1) * kBinStride) : kStatusInclusive;
  }

  // ---- scatter into shared memory in tile-sorted order (needs tile_prefix + warp offsets only)
#pragma unroll
  for (int i = 0; i < kItems; ++i) {
    const int local = wbase + 32 * i;
    if (kFull || local < tile_items) {
      const uint32_t bin = key_digit(key[i], d);
      const uint32_t pv = pay[i];
      const uint32_t r = (i & 1) ? (rank2[i >> 1] >> 16) : (rank2[i >> 1] & 0xFFFFu);
      const uint32_t pos = myhist[bin] + r;
      s.keys[pos] = key[i];
      s.pay[pos] = pv;
    }
  }
  // ---- resolve the look-back
  auto resolve = [&](const int b, uint32_t (&lk)[kLook]) {
    uint32_t excl = 0;
    if (tile > 0) {
      int64_t back = tile - 1;  // tile lk[0] belongs to
      uint32_t spins = 0;
      bool done = false;
      while (!done) {
        int j = 0;
#pragma unroll
        for (; j < kLook; ++j) {
          const uint32_t v = lk[j];
          const uint32_t status = v & ~kValueMask;
          if (status == 0) break;
          excl += v & kValueMask;
          if (status == kStatusInclusive) {
            done = true;
            break;
          }
        }
        if (done) break;
        back -= j;
        if (++spins > (1u << 22)) {  // never hang the GPU: flag and bail out
          atomicExch(p.err, 1u);
          break;
        }
        if (j == 0) __nanosleep(40);  // the predecessor has not published yet: do not burn issue slots
#pragma unroll
        for (int jj = 0; jj < kLook; ++jj)
          lk[jj] = back - jj >= 0 ? ld_volatile_u32(state + b - (tile - back + jj) * kBinStride) : kStatusInclusive;
      }
      st_volatile_u32(state + b, kStatusInclusive | ((excl + s.tile_count[b]) & kValueMask));
    }
    s.bin_delta[b] = base_hist[b] + excl - s.tile_prefix[b];
  };
  if (t < 256) resolve(t, look);
  __syncthreads();

  // ---- digit-contiguous global stores
  uint64_t* keys_dst = p.keys[info.first ? 0 : (info.src ^ 1)];
  uint32_t* keys32_dst = reinterpret_cast<uint32_t*>(keys_dst);
  uint32_t* pays_dst = p.pays[info.first ? 0 : (info.src ^ 1)];
#pragma unroll 8
  for (int k = 0; k < kItems; ++k) {
    const int j = k * kThreads + t;
    if (kFull || j < tile_items) {
      const KeyT kk = s.keys[j];
      const uint32_t pv = s.pay[j];
      const uint32_t bin = key_digit(kk, d);
      const uint32_t pos = uint32_t(j) + s.bin_delta[bin];
      if (info.last) {
        if (kPairs && info.hw) {
          // high word first over (key, payload) pairs: the position travelled through the passes (the repair looks
          // the lower key word up by it); the payload is fetched now and the position kept beside the output
          p.out_idx[pos] = uint64_t(p.pay_in[pv]);
          pays_dst[pos] = pv;
        } else {
          p.out_idx[pos] = p.gather ? p.gather[pv] : (uint64_t(pv) + p.index_base);
        }
        if (!kNarrow && p.out_keys) p.out_keys[pos] = kk;
        if (kNarrow && info.hw) keys32_dst[pos] = uint32_t(kk);   // the sorted upper words: the tie repair reads them
      } else {
        if (kNarrow) keys32_dst[pos] = uint32_t(kk);
        else if (info.out_narrow) keys32_dst[pos] = uint32_t(uint64_t(kk) >> (p.n_digits <= 4 ? 0 : 32));
        else keys_dst[pos] = kk;
        pays_dst[pos] = pv;
      }
    }
  }
}

// kLoop (round 2 of a failed high-word-first attempt, normally idle): a grid of a few CTAs per SM takes tiles until none
// is left — tickets are handed out in order, so every tile a look-back waits for is already running — and the idle case
// costs a few hundred CTAs that exit at once instead of one per tile
template <bool kNarrow, bool kLoop, bool kPairs = false>
__global__ void __launch_bounds__(kThreads, kNarrow ? B200_SORT_NARROW_MINBLOCKS : kMinBlocks) sort_pass_kernel(SortParams p, int d) {
  const PassInfo info = p.info[d];
  if (info.skip) return;
  if (kLoop && p.round2 && p.hws->fail == 0) return;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  using Smem = PassSmem<typename std::conditional<kNarrow, uint32_t, uint64_t>::type>;
  Smem& s = *reinterpret_cast<Smem*>(smem_raw);
  using KeyT = typename std::conditional<kNarrow, uint32_t, uint64_t>::type;
  const int t = threadIdx.x;
  for (;;) {
    if (t == 0) {
      const uint32_t tk = atomicAdd(p.counters + d, 1u);
      s.tile = tk;
      if (B200_SORT_TMA && !kLoop && !info.first && (int64_t(tk) + 1) * kTile <= p.n) {
        // a full tile of a later pass: both ranges go to the copy engine now
        mbar_init(&s.mbar, 1);
        mbar_expect_tx(&s.mbar, uint32_t(kTile * (sizeof(KeyT) + 4)));
        bulk_load(s.keys, reinterpret_cast<const KeyT*>(p.keys[info.src]) + size_t(tk) * kTile, uint32_t(kTile * sizeof(KeyT)), &s.mbar);
        bulk_load(s.pay, p.pays[info.src] + size_t(tk) * kTile, uint32_t(kTile * 4), &s.mbar);
      }
    }
    for (int i = t; i < kWarps * kBinStride / 4; i += kThreads) reinterpret_cast<uint4*>(&s.whist[0][0])[i] = make_uint4(0, 0, 0, 0);
    if (B200_SORT_MATCH_ATOMIC && !B200_SORT_TMA) {
      for (int i = t; i < kWarps * 256 / 4; i += kThreads) reinterpret_cast<uint4*>(&s.pay[0])[i] = make_uint4(0, 0, 0, 0);
    }
    __syncthreads();
    const int64_t tile = s.tile;
    if (kLoop && tile >= p.n_tiles) return;
    const int tile_items = int(min(int64_t(kTile), p.n - tile * kTile));
    if (tile_items == kTile) sort_pass_body<true, kNarrow, kPairs>(p, d, info, s, tile, tile_items);
    else sort_pass_body<false, kNarrow, kPairs>(p, d, info, s, tile, tile_items);
    if (!kLoop) return;
    __syncthreads();   // the staging arrays and the ticket are about to be reused
  }
}


// ---- NaN / null rows without a class pass ----------------------------------------------------------------------------
// Null and NaN rows carry the all-ones key through every pass like genuine maxima do, so a plain stable LSD sort leaves
// exactly the rows whose key is all ones at the very end, in row order.  Their number is counted by the histogram
// kernel; the three kernels below split that tail stably into [genuine maxima | NaNs | nulls] (vector_sort.cc:269-302).
// A column with 1 % nulls re-orders 1 % of the rows instead of ranking every row of the last pass over 258 bins with a
// ninth ballot and a validity gather (1.02 ms against 0.70 ms for a plain pass at 100 M rows, round 1).
constexpr int kTailThreads = 256;
constexpr int kTailMaxBlocks = 1024;
constexpr int kTailClasses = 4;   // other keys (high word first only) | genuine maxima | NaNs | nulls

struct TailArgs {
  SortParams p;
  uint64_t* idx;      // the sorted indices (out_idx)
  uint32_t* counts;   // [blocks][4] class counts, then exclusive positions
};

// the pass that wrote the indices (the plan lives in device memory)
__device__ __forceinline__ PassInfo last_pass(const SortParams& p) {
  PassInfo pi = p.info[0];
  for (int d = 0; d < kDigits; ++d)
    if (!p.info[d].skip && p.info[d].last) pi = p.info[d];
  return pi;
}
// scratch of the tail kernels: the key buffer the last pass read (free since)
__device__ __forceinline__ uint64_t* tail_tmp(const SortParams& p) {
  const PassInfo pi = last_pass(p);
  return p.keys[pi.first ? 1 : pi.src];
}

__device__ __forceinline__ uint32_t tail_class(const SortParams& p, uint64_t out_index) {
  const int64_t row = int64_t(out_index - p.index_base);
  if (!bitmap_get(p.vv, row)) return 3;
  uint32_t cls;
  const uint64_t k = transform_key(p, load_raw(p, row), &cls);
  if (cls) return 2;
  return k == p.wmask ? 1 : 0;   // 0: only in the run of the all-ones upper word of a high-word-first sort
}

// does this launch have anything to do?  (round 2 runs only after a failed high-word-first attempt)
__device__ __forceinline__ bool tail_active(const TailArgs& a) { return !(a.p.round2 && a.p.hws->fail == 0); }
// rows of the tail: the keys that are all ones — or, after the passes of a high-word-first plan, the whole run of the
// all-ones upper word
__device__ __forceinline__ int64_t tail_rows(const TailArgs& a) {
  const bool top = !a.p.round2 && a.p.hws && a.p.hws->hw != 0;
  return top ? a.p.n - a.p.hws->top_start : int64_t(*a.p.tail_count);
}

__device__ __forceinline__ void tail_range(const TailArgs& a, int64_t* begin, int64_t* end, int64_t* first) {
  const int64_t T = tail_rows(a);
  int64_t chunk = (T + gridDim.x - 1) / gridDim.x;
  chunk = (chunk + kTailThreads - 1) / kTailThreads * kTailThreads;
  *first = a.p.n - T;
  *begin = min(T, int64_t(blockIdx.x) * chunk);
  *end = min(T, *begin + chunk);
}

__global__ void __launch_bounds__(kTailThreads) tail_count_kernel(const __grid_constant__ TailArgs a) {
  if (!tail_active(a)) return;
  __shared__ uint32_t c[kTailClasses];
  if (threadIdx.x < kTailClasses) c[threadIdx.x] = 0;
  __syncthreads();
  int64_t begin, end, first;
  tail_range(a, &begin, &end, &first);
  for (int64_t i = begin + threadIdx.x; i < end; i += kTailThreads) atomicAdd(&c[tail_class(a.p, a.idx[first + i])], 1u);
  __syncthreads();
  if (threadIdx.x < kTailClasses) a.counts[blockIdx.x * kTailClasses + threadIdx.x] = c[threadIdx.x];
}

// one CTA: counts[b][c] -> where block b's rows of class c start inside the tail; the class totals go to hws
__global__ void __launch_bounds__(kTailMaxBlocks) tail_scan_kernel(const __grid_constant__ TailArgs a, int blocks) {
  if (!tail_active(a)) return;
  uint32_t* counts = a.counts;
  __shared__ uint32_t warp_sums[32];
  __shared__ uint32_t class_start;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  if (t == 0) class_start = 0;
  __syncthreads();
  for (int c = 0; c < kTailClasses; ++c) {
    const uint32_t v = t < blocks ? counts[t * kTailClasses + c] : 0;
    uint32_t incl = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += o;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    uint32_t wpre = 0, total = 0;
    for (int w = 0; w < kTailMaxBlocks / 32; ++w) {
      if (w < warp) wpre += warp_sums[w];
      total += warp_sums[w];
    }
    const uint32_t start = class_start;
    if (t < blocks) counts[t * kTailClasses + c] = start + wpre + incl - v;
    __syncthreads();
    if (t == 0) {
      class_start = start + total;
      if (a.p.hws) a.p.hws->class_total[c] = total;
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(kTailThreads) tail_scatter_kernel(const __grid_constant__ TailArgs a) {
  if (!tail_active(a)) return;
  __shared__ uint32_t run[kTailClasses];                        // next position of every class for this block
  __shared__ uint32_t wcnt[kTailThreads / 32][kTailClasses];
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const unsigned lt = lanemask_lt();
  uint64_t* tmp = tail_tmp(a.p);
  if (t < kTailClasses) run[t] = a.counts[blockIdx.x * kTailClasses + t];
  int64_t begin, end, first;
  tail_range(a, &begin, &end, &first);
  for (int64_t base = begin; base < end; base += kTailThreads) {
    const int64_t i = base + t;
    const bool in = i < end;
    const uint64_t v = in ? a.idx[first + i] : 0;
    const uint32_t c = in ? tail_class(a.p, v) : kTailClasses;
    uint32_t rank = 0;
#pragma unroll
    for (int q = 0; q < kTailClasses; ++q) {
      const unsigned m = __ballot_sync(0xffffffffu, c == q);
      if (c == q) rank = __popc(m & lt);
      if (lane == 0) wcnt[warp][q] = __popc(m);
    }
    __syncthreads();
    if (in) {
      uint32_t pre = run[c];
      for (int w = 0; w < warp; ++w) pre += wcnt[w][c];
      tmp[pre + rank] = v;
    }
    __syncthreads();
    if (t < kTailClasses) {
      uint32_t add = 0;
      for (int w = 0; w < kTailThreads / 32; ++w) add += wcnt[w][t];
      run[t] += add;
    }
    __syncthreads();
  }
}

__global__ void tail_copy_kernel(const __grid_constant__ TailArgs a) {
  if (!tail_active(a)) return;
  const int64_t T = tail_rows(a);
  uint64_t* dst = a.idx + (a.p.n - T);
  const uint64_t* tmp = tail_tmp(a.p);
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < T; i += int64_t(gridDim.x) * blockDim.x) dst[i] = tmp[i];
}

// ---- high word first: the repair of the ties (see HwState) ------------------------------------------------------------

// the sorted upper words the last pass of a high-word-first plan left behind
__device__ __forceinline__ const uint32_t* hw_words(const SortParams& p) {
  const PassInfo pi = last_pass(p);
  return reinterpret_cast<const uint32_t*>(p.keys[pi.first ? 0 : (pi.src ^ 1)]);
}

// where the run of the all-ones upper word starts (the words are sorted: a binary search by one thread)
__global__ void hw_top_kernel(SortParams p) {
  if (p.hws->hw == 0) return;
  const uint32_t* hi = hw_words(p);
  int64_t lo = 0, up = p.n;   // first position whose word is all ones
  while (lo < up) {
    const int64_t mid = (lo + up) >> 1;
    if (hi[mid] == 0xFFFFFFFFu) up = mid;
    else lo = mid + 1;
  }
  p.hws->top_start = lo;
}

// where the repair may look: with nulls / NaNs in play the run of the all-ones word was split by the tail kernels and only
// its first part (keys that are not all ones) is a run to repair
__device__ __forceinline__ int64_t hw_limit(const SortParams& p) {
  return p.tail_fix ? p.hws->top_start + int64_t(p.hws->class_total[0]) : p.n;
}
// the list of run starts: the payload buffer the last pass did not write is free by now
__device__ __forceinline__ uint32_t* hw_run_list(const SortParams& p) {
  const PassInfo pi = last_pass(p);
  return p.pays[pi.first ? 1 : pi.src];
}

// repair, step 1 (streaming): the positions where a run of equal upper words starts, appended to a list.  A CTA takes
// 4096 positions at a time, 16 consecutive ones per thread, and claims room in the list with ONE atomic per tile (a
// warp-level claim per 32 positions serialised a million atomics on one address: 0.6 ms).  (One kernel that repaired
// a run where it met it ran at a fifth of the memory bandwidth: a warp in which one lane had found a run waited for
// that lane's three dependent gathers, every third warp iteration.)
constexpr int kHwRunsPerThread = 16;
__global__ void __launch_bounds__(256) hw_runs_kernel(SortParams p) {
  if (p.hws->hw == 0) return;
  __shared__ uint32_t warp_sums[8];
  __shared__ uint32_t tile_base;
  const uint32_t* hi = hw_words(p);
  uint32_t* list = hw_run_list(p);
  const int64_t limit = hw_limit(p);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  constexpr int kTilePositions = 256 * kHwRunsPerThread;
  for (int64_t base = int64_t(blockIdx.x) * kTilePositions; base < limit; base += int64_t(gridDim.x) * kTilePositions) {
    const int64_t i0 = base + int64_t(t) * kHwRunsPerThread;
    // words i0 - 1 .. i0 + 16 (the buffer is 256-byte aligned and i0 a multiple of 16: four aligned 16-byte loads)
    uint32_t w[kHwRunsPerThread + 2];
    if (i0 + kHwRunsPerThread <= limit) {
#pragma unroll
      for (int q = 0; q < kHwRunsPerThread / 4; ++q) {
        const uint4 v = reinterpret_cast<const uint4*>(hi + i0)[q];
        w[1 + 4 * q] = v.x;
        w[2 + 4 * q] = v.y;
        w[3 + 4 * q] = v.z;
        w[4 + 4 * q] = v.w;
      }
    } else {
#pragma unroll
      for (int q = 0; q < kHwRunsPerThread; ++q) w[1 + q] = i0 + q < limit ? hi[i0 + q] : 0u;
    }
    w[0] = i0 > 0 && i0 - 1 < limit ? hi[i0 - 1] : 0u;
    w[kHwRunsPerThread + 1] = i0 + kHwRunsPerThread < limit ? hi[i0 + kHwRunsPerThread] : 0u;
    uint32_t starts = 0;
#pragma unroll
    for (int q = 0; q < kHwRunsPerThread; ++q) {
      const int64_t i = i0 + q;
      const bool start = i + 1 < limit && w[q + 2] == w[q + 1] && (i == 0 || w[q] != w[q + 1]);
      starts |= uint32_t(start) << q;
    }
    // exclusive scan of the threads' counts over the CTA
    const uint32_t c = __popc(starts);
    uint32_t incl = c;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += o;
    }
    if (lane == 31) warp_sums[warp] = incl;
    __syncthreads();
    uint32_t wpre = 0, total = 0;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (k < warp) wpre += warp_sums[k];
      total += warp_sums[k];
    }
    if (t == 0) tile_base = total ? atomicAdd(&p.hws->n_runs, total) : 0u;
    __syncthreads();
    uint32_t at = tile_base + wpre + incl - c;
    while (starts) {
      const int q = __ffs(starts) - 1;
      starts &= starts - 1;
      list[at++] = uint32_t(i0 + q);
    }
    __syncthreads();   // warp_sums / tile_base are rewritten by the next tile
  }
}

// repair, step 2: one thread per run puts it in (lower word, row) order; every thread is gathering, so the dependent
// loads (index -> row -> column value) of thousands of runs are in flight together
__global__ void __launch_bounds__(256) hw_fix_kernel(SortParams p) {
  if (p.hws->hw == 0) return;
  const uint32_t* hi = hw_words(p);
  const uint32_t* list = hw_run_list(p);
  const int64_t limit = hw_limit(p);
  const uint32_t n_runs = p.hws->n_runs;
  const PassInfo last = last_pass(p);
  const uint32_t* positions = p.pay_in ? p.pays[last.first ? 0 : (last.src ^ 1)] : nullptr;
  for (uint32_t r = blockIdx.x * blockDim.x + threadIdx.x; r < n_runs; r += gridDim.x * blockDim.x) {
    const int64_t i = list[r];
    const uint32_t w = hi[i];
    int len = 2;
    while (len <= kHwMaxRun && i + len < limit && hi[i + len] == w) ++len;
    if (len > kHwMaxRun) {   // the estimate of the plan kernel was wrong: the ordinary plan runs after all
      p.hws->fail = 1u;
      continue;
    }
    uint64_t idx[kHwMaxRun];
    uint32_t low[kHwMaxRun];
    for (int j = 0; j < len; ++j) {
      const uint64_t v = p.out_idx[i + j];
      uint32_t cls;
      // where the element came from: its row (column sorts: the output is the row number) or, for (key, payload)
      // pairs, the position the last pass left beside the output
      const int64_t from = positions ? int64_t(positions[i + j]) : int64_t(v - p.index_base);
      const uint32_t l = uint32_t(source_key(p, from, &cls));
      // stable insertion: the rows arrive in row order
      int q = j;
      while (q > 0 && low[q - 1] > l) {
        low[q] = low[q - 1];
        idx[q] = idx[q - 1];
        --q;
      }
      low[q] = l;
      idx[q] = v;
    }
    for (int j = 0; j < len; ++j) p.out_idx[i + j] = idx[j];
  }
}

// round 2 needs the look-back state of digits 4..7 clean again
__global__ void hw_rezero_kernel(SortParams p) {
  if (p.hws->fail == 0) return;
  uint4* s = reinterpret_cast<uint4*>(p.state + size_t(4) * p.n_tiles * kBinStride);
  const size_t words = size_t(4) * p.n_tiles * kBinStride / 4;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < words; i += size_t(gridDim.x) * blockDim.x) s[i] = make_uint4(0, 0, 0, 0);
}

__global__ void sort_keys_kernel(SortParams p, uint64_t* out, uint64_t* out_rows) {
  const int64_t row = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (row < p.n) {
    uint32_t cls;
    const uint64_t k = source_key(p, row, &cls);
    out[row] = cls == 2 ? kNullKeyOut : (cls == 1 ? kNanKeyOut : k);
    if (out_rows) out_rows[row] = (p.index_base + uint64_t(row)) | (uint64_t(cls) << 62);
  }
}

// ---- stable range partition (multi-GPU sample sort) ----------------------------------
// dest(row) = number of splitters <= key(row) (unsigned radix keys; NaN -> 2^64-2, null -> 2^64-1 as in
// sort_keys_kernel).  Rows of one destination come out contiguous AND in row order, so that after the
// all-to-all the concatenation of the runs a rank receives (source ranks in rank order) is in global row
// order and ONE stable sort by key finishes the job — the first version sorted locally before the exchange
// and again after it.  256 threads x 16 rows; every warp owns 512 consecutive rows.
constexpr int kSpMax = 16;
constexpr int kSpThreads = 256;
constexpr int kSpItems = 16;
constexpr int kSpTile = kSpThreads * kSpItems;

struct Splitters {
  uint64_t s[kSpMax - 1];
  int32_t n_parts;
};

// number of splitters <= key: the splitters ascend, so four halving steps replace the 15 compares
__device__ __forceinline__ uint32_t sp_dest(const Splitters& sp, uint64_t key) {
  const uint32_t ns = uint32_t(sp.n_parts - 1);
  uint32_t d = 0;
#pragma unroll
  for (uint32_t step = 8; step > 0; step >>= 1) {
    const uint32_t j = d + step;   // candidate count: valid if splitter j-1 exists and is <= key
    if (j <= ns && sp.s[j - 1] <= key) d = j;
  }
  return d;
}

__device__ __forceinline__ uint64_t sp_out_key(const SortParams& p, int64_t row, uint32_t* cls) {
  const uint64_t k = source_key(p, row, cls);
  return *cls == 2 ? kNullKeyOut : (*cls == 1 ? kNanKeyOut : k);
}

__global__ void __launch_bounds__(kSpThreads) sp_count_kernel(SortParams p, const __grid_constant__ Splitters sp, uint32_t* tile_counts) {
  __shared__ uint32_t cnt[kSpMax];
  const int lane = threadIdx.x & 31;
  if (threadIdx.x < kSpMax) cnt[threadIdx.x] = 0;
  __syncthreads();
  const int64_t base = int64_t(blockIdx.x) * kSpTile;
  // lane q (< 16) counts the warp's rows for destination q from four ballots per step: with 2..16 destinations
  // per-row shared atomics would all land on a handful of addresses and serialise
  uint32_t mine = 0;
#pragma unroll 4
  for (int i = 0; i < kSpItems; ++i) {
    const int64_t row = base + i * kSpThreads + threadIdx.x;
    const bool in = row < p.n;
    uint32_t d = 0;
    if (in) {
      uint32_t cls;
      d = sp_dest(sp, sp_out_key(p, row, &cls));
    }
    unsigned mq = __ballot_sync(0xffffffffu, in);
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const unsigned bal = __ballot_sync(0xffffffffu, in && ((d >> b) & 1u));
      mq &= ((lane >> b) & 1u) ? bal : ~bal;
    }
    mine += __popc(mq);
  }
  if (lane < kSpMax && mine) atomicAdd(&cnt[lane], mine);
  __syncthreads();
  if (threadIdx.x < kSpMax) tile_counts[size_t(blockIdx.x) * kSpMax + threadIdx.x] = cnt[threadIdx.x];
}

// one CTA per destination: exclusive scan of its per-tile counts; totals[q] = count.  The exclusive bases
// over destinations (totals[kSpMax + q]) are written by the last CTA to finish.
__global__ void __launch_bounds__(1024) sp_scan_kernel(uint32_t* tile_counts, int64_t n_tiles, int n_parts, int64_t* totals,
                                                       uint32_t* done_counter) {
  __shared__ uint32_t warp_sums[32];
  __shared__ uint32_t carry_s;
  __shared__ bool last_s;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int q = blockIdx.x;
  if (t == 0) carry_s = 0;
  __syncthreads();
  for (int64_t base = 0; base < n_tiles; base += 1024) {
    const int64_t tile = base + t;
    const uint32_t c = tile < n_tiles ? tile_counts[size_t(tile) * kSpMax + q] : 0;
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
    if (tile < n_tiles) tile_counts[size_t(tile) * kSpMax + q] = carry + wpre + incl - c;
    __syncthreads();
    if (t == 1023) carry_s = carry + wpre + incl;
    __syncthreads();
  }
  if (t == 0) {
    totals[q] = carry_s;
    __threadfence();
    last_s = atomicAdd(done_counter, 1u) == uint32_t(n_parts - 1);
  }
  __syncthreads();
  if (last_s && t == 0) {
    __threadfence();
    int64_t acc = 0;
    for (int j = 0; j < n_parts; ++j) {
      totals[kSpMax + j] = acc;
      acc += reinterpret_cast<volatile int64_t*>(totals)[j];
    }
    *done_counter = 0;   // ready for the next call (stream-ordered)
  }
}

// Where the scatter writes.  Local partition: every destination shares one buffer, run q starts at the
// device-side totals[kSpMax + q].  Fused exchange: keys[q] / rows[q] are rank q's receive buffers (peer
// mappings over NVLink; this rank's own for q == rank) and base[q] where this rank's run starts in them.
struct PushDst {
  uint64_t* keys[kSpMax];
  void* rows[kSpMax];
  int64_t base[kSpMax];
  const int64_t* totals;  // nullable
  int32_t rows32;
};

// The tile's rows are first grouped by destination in shared memory, then each destination's run (tile rows /
// n_parts long) leaves as one contiguous stream of full-width warp stores.  Peer stores are not combined by the
// local L2, so writing straight from the ranking registers sent NVLink one 8..32-byte fragment per (warp store,
// destination): 150 GB/s at 8 GPUs.
__global__ void __launch_bounds__(kSpThreads) sp_scatter_kernel(SortParams p, const __grid_constant__ Splitters sp,
                                                                const uint32_t* tile_offsets,
                                                                const __grid_constant__ PushDst dst) {
  __shared__ uint64_t skeys[kSpTile];
  __shared__ uint16_t srow[kSpTile];          // row inside the tile | class << 12
  __shared__ uint32_t wtot[kSpThreads / 32][kSpMax];
  __shared__ uint32_t run_start[kSpMax + 1];  // tile-local start of every destination's run; [n_parts] = rows in the tile
  __shared__ uint64_t gbase[kSpMax];          // where the tile's run starts in destination q's buffer
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const unsigned lt = lanemask_lt();
  const int64_t tile_row0 = int64_t(blockIdx.x) * kSpTile;
  const int64_t wrow0 = tile_row0 + warp * (kSpItems * 32);
  uint64_t key[kSpItems];
  uint32_t meta[kSpItems];   // dest | class << 4 | rank among the warp's rows for that destination << 8
  uint32_t cnt = 0;          // lane q (< 16): rows of this warp seen so far for destination q
#pragma unroll
  for (int i = 0; i < kSpItems; ++i) {
    const int64_t row = wrow0 + i * 32 + lane;
    const bool in = row < p.n;
    uint32_t cls = 0, d = 0;
    key[i] = 0;
    if (in) {
      key[i] = sp_out_key(p, row, &cls);
      d = sp_dest(sp, key[i]);
    }
    const unsigned inmask = __ballot_sync(0xffffffffu, in);
    unsigned m = inmask, mq = inmask;
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const unsigned bal = __ballot_sync(0xffffffffu, in && ((d >> b) & 1u));
      m &= ((d >> b) & 1u) ? bal : ~bal;
      mq &= ((lane >> b) & 1u) ? bal : ~bal;
    }
    const uint32_t before = __shfl_sync(0xffffffffu, cnt, int(d));
    meta[i] = d | (cls << 4) | ((before + __popc(m & lt)) << 8);
    cnt += __popc(mq);
  }
  if (lane < kSpMax) wtot[warp][lane] = cnt;
  __syncthreads();
  if (warp == 0) {
    // lane q: the tile's rows for destination q, their exclusive prefix over q, and every warp's start
    uint32_t tq = 0;
    if (lane < kSpMax) {
#pragma unroll
      for (int w = 0; w < kSpThreads / 32; ++w) tq += wtot[w][lane];
    }
    uint32_t incl = tq;
#pragma unroll
    for (int sft = 1; sft < kSpMax; sft <<= 1) {
      const uint32_t o = __shfl_up_sync(0xffffffffu, incl, sft);
      if (lane >= sft) incl += o;
    }
    if (lane < kSpMax) {
      uint32_t acc = incl - tq;
      run_start[lane] = acc;
      if (lane == kSpMax - 1) run_start[kSpMax] = incl;
#pragma unroll
      for (int w = 0; w < kSpThreads / 32; ++w) {
        const uint32_t c = wtot[w][lane];
        wtot[w][lane] = acc;
        acc += c;
      }
      gbase[lane] = lane < sp.n_parts
                        ? uint64_t(dst.base[lane]) + (dst.totals ? uint64_t(dst.totals[kSpMax + lane]) : 0) +
                              tile_offsets[size_t(blockIdx.x) * kSpMax + lane]
                        : 0;
    }
  }
  __syncthreads();
  const uint32_t off = lane < kSpMax ? wtot[warp][lane] : 0;
#pragma unroll
  for (int i = 0; i < kSpItems; ++i) {
    const uint32_t d = meta[i] & 15u;
    const uint32_t lpos = __shfl_sync(0xffffffffu, off, int(d)) + (meta[i] >> 8);
    if (wrow0 + i * 32 + lane < p.n) {
      skeys[lpos] = key[i];
      srow[lpos] = uint16_t(uint32_t(warp * (kSpItems * 32) + i * 32 + lane) | (((meta[i] >> 4) & 3u) << 12));
    }
  }
  __syncthreads();
  const uint32_t tile_rows = run_start[kSpMax];
#pragma unroll 4
  for (int k = 0; k < kSpItems; ++k) {
    const uint32_t j = uint32_t(k * kSpThreads + t);
    if (j < tile_rows) {
      uint32_t q = 0;
#pragma unroll
      for (uint32_t step = 8; step > 0; step >>= 1)   // last destination whose run starts at or before j
        if (q + step < uint32_t(sp.n_parts) && run_start[q + step] <= j) q += step;
      const uint64_t pos = gbase[q] + (j - run_start[q]);
      const uint32_t sr = srow[j];
      const uint64_t row = p.index_base + uint64_t(tile_row0 + (sr & 4095u));
      const uint32_t cls = sr >> 12;
      dst.keys[q][pos] = skeys[j];
      if (dst.rows32) reinterpret_cast<uint32_t*>(dst.rows[q])[pos] = uint32_t(row) | (cls << 30);
      else reinterpret_cast<uint64_t*>(dst.rows[q])[pos] = row | (uint64_t(cls) << 62);
    }
  }
}

// ---- chunked input: class keys for the final stable 3-way split, and the class-bit strip
// sorted (key, row | class << 62) pairs -> 0 for keys below the NaN key, else 1 + class: one more stable
// pass on this 2-bit key puts [values | genuine maxima | NaNs | nulls] in place (vector_sort.cc:269-302)
__global__ void chunk_class_keys_kernel(const uint64_t* keys, const uint64_t* rows, int64_t n, uint64_t* out) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) out[i] = keys[i] >= kNanKeyOut ? 1 + (rows[i] >> 62) : 0;
}
__global__ void chunk_strip_kernel(uint64_t* rows, int64_t n) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) rows[i] &= (uint64_t(1) << 62) - 1;
}

// ---- several sort keys (SortIndices on a RecordBatch / Table): LSD over the keys, least significant first.
// Each key is one stable sort of (value key, row | class << 30) pairs: the rows arrive in the order the less
// significant keys left them (perm; nullptr = identity), nulls and NaNs share the all-ones value key so the
// stable passes keep that order, and one more stable pass on the 2-bit class moves NaNs, then nulls, behind the
// values (MultipleKeyComparator::CompareType, vector_sort.cc:1098-1170).
__global__ void mk_keys_kernel(SortParams p, const uint64_t* perm, uint64_t* keys, uint32_t* pay) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < p.n) {
    const int64_t row = perm ? int64_t(perm[i]) : i;
    uint32_t cls;
    keys[i] = source_key(p, row, &cls);
    pay[i] = uint32_t(row) | (cls << 30);
  }
}
__global__ void mk_class_kernel(const uint64_t* rows, int64_t n, uint64_t* ckeys, uint32_t* pay) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) {
    ckeys[i] = rows[i] >> 30;
    pay[i] = uint32_t(rows[i]) & 0x3fffffffu;
  }
}

// chunked key columns: the key's (value key, class) of every row in row order first, then gathered through
// the permutation like mk_keys_kernel does from a single column
__global__ void mk_flat_kernel(SortParams p, int64_t off, uint64_t* flat_keys, uint8_t* flat_cls) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < p.n) {
    uint32_t cls;
    flat_keys[off + i] = source_key(p, i, &cls);
    flat_cls[off + i] = uint8_t(cls);
  }
}
__global__ void mk_gather_flat_kernel(const uint64_t* flat_keys, const uint8_t* flat_cls, const uint64_t* perm, int64_t n,
                                      uint64_t* keys, uint32_t* pay) {
  const int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i < n) {
    const uint64_t row = perm[i];
    keys[i] = flat_keys[row];
    pay[i] = uint32_t(row) | (uint32_t(flat_cls[row]) << 30);
  }
}

struct Layout {
  size_t hist, info, info2, hw, counters, err, tail, state, zero_bytes, keys0, keys1, pays0, pays1, total;
};

Layout layout_for(int64_t n) {
  Layout L;
  const size_t n_tiles = size_t((n + kTile - 1) / kTile);
  size_t off = 0;
  L.hist = off; off += b200_align_up(size_t(kDigits) * kBinStride * 4, 256);
  L.info = off; off += 256;
  L.info2 = off; off += 256;
  L.hw = off; off += 256;      // HwState, then round 2's tile tickets (+64) and tail count (+128)
  L.counters = off; off += 128;
  L.err = off; off += 128;
  L.tail = off; off += size_t(kTailMaxBlocks) * 16;
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
  p->width = bits / 8;
  switch (values->type) {
    case B200_FLOAT: case B200_DOUBLE: p->tkind = kFloatKey; break;
    case B200_INT8: case B200_INT16: case B200_INT32: case B200_INT64: p->tkind = kSignedKey; break;
    default: p->tkind = kUnsignedKey; break;
  }
  p->sign = uint64_t(1) << (bits - 1);
  p->wmask = bits == 64 ? ~uint64_t(0) : ((uint64_t(1) << bits) - 1);
  p->inv = order ? p->wmask : 0;
  p->nan_above = bits == 64 ? 0x7FF0000000000000ull : 0x7F800000ull;
  p->n_digits = bits / 8;
  p->cls_digit = (has_nulls || p->tkind == kFloatKey) ? bits / 8 - 1 : -1;
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
  // the look-back's bail-out flag lives in the context: b200_ctx_synchronize reports it (a preempted or time-sliced
  // GPU must not turn into a silently wrong permutation)
  p.err = ctx->async_err;
  p.state = reinterpret_cast<uint32_t*>(base + L.state);
  p.keys[0] = reinterpret_cast<uint64_t*>(base + L.keys0);
  p.keys[1] = reinterpret_cast<uint64_t*>(base + L.keys1);
  p.pays[0] = reinterpret_cast<uint32_t*>(base + L.pays0);
  p.pays[1] = reinterpret_cast<uint32_t*>(base + L.pays1);
  // only the digits that can vary need look-back state
  const size_t zero_bytes = L.state + b200_align_up(size_t(p.n_digits) * size_t(p.n_tiles) * kBinStride * 4, 256);
  B200_CUDA(ctx, cudaMemsetAsync(base, 0, std::min(zero_bytes, L.zero_bytes), ctx->stream));
  const int hist_blocks = int(std::min<int64_t>((p.n + kHistTile - 1) / kHistTile, int64_t(ctx->num_sms) * 8));
  const int smem_wide = int(sizeof(PassSmem<uint64_t>)), smem_narrow = int(sizeof(PassSmem<uint32_t>));
  if (b200_ctx_once(ctx, 0)) {
    B200_CUDA(ctx, cudaFuncSetAttribute(sort_pass_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_wide));
    B200_CUDA(ctx, cudaFuncSetAttribute(sort_pass_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_narrow));
    B200_CUDA(ctx, cudaFuncSetAttribute(sort_pass_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_wide));
    B200_CUDA(ctx, cudaFuncSetAttribute(sort_pass_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_narrow));
    B200_CUDA(ctx, cudaFuncSetAttribute(sort_pass_kernel<true, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_narrow));
  }
  // the caller wants the sorted 64-bit keys back: every pass carries them whole
  p.narrow = p.out_keys ? 0 : 1;
  // a column that can hold nulls or NaNs: plain passes, then the tail fix-up (layout slot `err` is free: the bail-out
  // flag lives in the context)
  p.tail_fix = p.cls_digit >= 0 ? 1 : 0;
  p.tail_count = reinterpret_cast<uint32_t*>(base + L.err);
  // high word first (HwState): a 64-bit column sorted for its indices alone; the plan kernel decides on the device
  p.info2 = reinterpret_cast<PassInfo*>(base + L.info2);
  p.hws = reinterpret_cast<HwState*>(base + L.hw);
  p.round2 = 0;
  p.hw_allowed = 0;
  // ... or 64-bit ready-made keys with a 32-bit payload (the second key of a multi-key sort, the received pairs of the
  // multi-GPU sample sort)
  const bool hw_column = !p.keys_in && p.width == 8;
  const bool hw_pairs = p.keys_in && p.pay_in && p.n_digits == kDigits;
  // (left to the estimate only from 2^18 rows on: the attempt costs a dozen extra launches, more than a small sort takes)
  if ((hw_column || hw_pairs) && p.narrow && !p.gather && ctx->sort_high_word_first >= 0 &&
      (ctx->sort_high_word_first > 0 || p.n >= (int64_t(1) << 18)))
    p.hw_allowed = ctx->sort_high_word_first > 0 ? 2 : 1;
  if ((p.keys_in != nullptr || p.width == 8) && p.n_digits == kDigits)
    B200_TIMED(ctx, "sort_hist", (sort_hist_kernel<true><<<hist_blocks, kHistThreads, 0, ctx->stream>>>(p)));
  else
    B200_TIMED(ctx, "sort_hist", (sort_hist_kernel<false><<<hist_blocks, kHistThreads, 0, ctx->stream>>>(p)));
  sort_scan_kernel<<<1, kHistThreads, 0, ctx->stream>>>(p);
  B200_LAUNCHED(ctx);
  const int tail_blocks = std::min(kTailMaxBlocks, ctx->num_sms * 4);
  auto passes = [&](const SortParams& q) -> int {
    for (int d = 0; d < q.n_digits; ++d) {
      const bool narrow = q.narrow && (d >= 4 || q.n_digits <= 4);
      const unsigned grid = unsigned(q.n_tiles);
      if (q.round2) {   // normally every launch exits at once
        const unsigned loop_grid = unsigned(std::min<int64_t>(q.n_tiles, int64_t(ctx->num_sms) * 3));
        if (narrow)
          B200_TIMED(ctx, "sort_pass_round2", (sort_pass_kernel<true, true><<<loop_grid, kThreads, smem_narrow, ctx->stream>>>(q, d)));
        else
          B200_TIMED(ctx, "sort_pass_round2", (sort_pass_kernel<false, true><<<loop_grid, kThreads, smem_wide, ctx->stream>>>(q, d)));
      } else if (!narrow && q.hw_allowed && B200_SORT_LOOP_LOW) {
        // digits 0..3 of a sort that may take the high-word-first plan (which skips them): the looping form, a few
        // hundred CTAs instead of one per tile, so that a skipped pass costs 4 us instead of 17
        const unsigned loop_grid = unsigned(std::min<int64_t>(q.n_tiles, int64_t(ctx->num_sms) * 3));
        B200_TIMED(ctx, "sort_pass", (sort_pass_kernel<false, true><<<loop_grid, kThreads, smem_wide, ctx->stream>>>(q, d)));
      } else if (narrow && q.hw_allowed && q.pay_in) {
        B200_TIMED(ctx, "sort_pass", (sort_pass_kernel<true, false, true><<<grid, kThreads, smem_narrow, ctx->stream>>>(q, d)));
      } else if (narrow) {
        B200_TIMED(ctx, "sort_pass", (sort_pass_kernel<true, false><<<grid, kThreads, smem_narrow, ctx->stream>>>(q, d)));
      } else {
        B200_TIMED(ctx, "sort_pass", (sort_pass_kernel<false, false><<<grid, kThreads, smem_wide, ctx->stream>>>(q, d)));
      }
    }
    return B200_OK;
  };
  auto tail = [&](const SortParams& q) -> int {
    TailArgs ta;
    ta.p = q;
    ta.idx = q.out_idx;
    ta.counts = reinterpret_cast<uint32_t*>(base + L.tail);
    B200_TIMED(ctx, "sort_tail", (tail_count_kernel<<<tail_blocks, kTailThreads, 0, ctx->stream>>>(ta)));
    tail_scan_kernel<<<1, kTailMaxBlocks, 0, ctx->stream>>>(ta, tail_blocks);
    B200_LAUNCHED(ctx);
    B200_TIMED(ctx, "sort_tail", (tail_scatter_kernel<<<tail_blocks, kTailThreads, 0, ctx->stream>>>(ta)));
    B200_TIMED(ctx, "sort_tail", (tail_copy_kernel<<<tail_blocks, kTailThreads, 0, ctx->stream>>>(ta)));
    return B200_OK;
  };
  B200_RETURN_NOT_OK(passes(p));
  if (p.hw_allowed && p.tail_fix) {
    hw_top_kernel<<<1, 1, 0, ctx->stream>>>(p);
    B200_LAUNCHED(ctx);
  }
  if (p.tail_fix) B200_RETURN_NOT_OK(tail(p));
  if (p.hw_allowed) {
    const unsigned fix_grid = unsigned(std::min<int64_t>((p.n + 255) / 256, int64_t(ctx->num_sms) * 16));
    B200_TIMED(ctx, "sort_tie_repair", (hw_runs_kernel<<<fix_grid, 256, 0, ctx->stream>>>(p)));
    B200_TIMED(ctx, "sort_tie_repair", (hw_fix_kernel<<<fix_grid, 256, 0, ctx->stream>>>(p)));
    // round 2: the ordinary plan, every launch gated on the failure flag of the repair (normally they all exit at once)
    SortParams q = p;
    q.round2 = 1;
    q.info = p.info2;
    q.counters = reinterpret_cast<uint32_t*>(base + L.hw + 64);
    q.tail_count = reinterpret_cast<uint32_t*>(base + L.hw + 128);
    hw_rezero_kernel<<<ctx->num_sms * 4, 256, 0, ctx->stream>>>(q);
    B200_LAUNCHED(ctx);
    B200_RETURN_NOT_OK(passes(q));
    if (q.tail_fix) B200_RETURN_NOT_OK(tail(q));
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

int b200_sort_indices_plan(b200_ctx* ctx, const void* scratch, int* out_high_word_first, int* out_fell_back) {
  B200_ENTER(ctx);
  if (!ctx || !scratch) return B200_INVALID;
  // the plan of the last b200_sort_indices call that used `scratch` lives at a fixed place inside it
  const Layout L = layout_for(0);
  const char* base = reinterpret_cast<const char*>(b200_align_up(reinterpret_cast<uintptr_t>(scratch), 256));
  HwState h;
  B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  B200_CUDA(ctx, cudaMemcpy(&h, base + L.hw, sizeof(h), cudaMemcpyDeviceToHost));
  if (out_high_word_first) *out_high_word_first = h.hw ? 1 : 0;
  if (out_fell_back) *out_fell_back = h.fail ? 1 : 0;
  return B200_OK;
}

int b200_sort_pairs_scratch_bytes(int64_t length, size_t* out_bytes) {
  return b200_sort_indices_scratch_bytes(length, out_bytes);
}

int b200_sort_indices(b200_ctx* ctx, const b200_col* values, int order, uint64_t index_base,
                      uint64_t* out_indices, void* scratch, size_t scratch_bytes) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  SortParams p;
  memset(&p, 0, sizeof(p));
  B200_RETURN_NOT_OK(fill_column_source(ctx, values, order, &p));
  if (p.n > 0 && !out_indices) return b200_fail(ctx, B200_INVALID, "sort_indices: null output");
  p.index_base = index_base;
  p.out_idx = out_indices;
  return run_sort(ctx, p, scratch, scratch_bytes);
}

int b200_sort_keys_rows(b200_ctx* ctx, const b200_col* values, int order, uint64_t index_base, uint64_t* out_keys,
                        uint64_t* out_rows) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  SortParams p;
  memset(&p, 0, sizeof(p));
  B200_RETURN_NOT_OK(fill_column_source(ctx, values, order, &p));
  if (p.n == 0) return B200_OK;
  if (!out_keys) return b200_fail(ctx, B200_INVALID, "sort_keys: null output");
  if (out_rows && index_base + uint64_t(p.n) > (uint64_t(1) << 62))
    return b200_fail(ctx, B200_INDEX_ERROR, "sort_keys: row numbers beyond 2^62");
  p.index_base = index_base;
  sort_keys_kernel<<<(unsigned)((p.n + 255) / 256), 256, 0, ctx->stream>>>(p, out_keys, out_rows);
  B200_LAUNCHED(ctx);
  return B200_OK;
}

int b200_sort_indices_chunked_scratch_bytes(int64_t total_length, size_t* out_bytes) {
  if (!out_bytes || total_length < 0) return B200_INVALID;
  // keys, rows, sorted keys, sorted rows (8 B each) + the pair sort's own scratch
  *out_bytes = 4 * b200_align_up(size_t(total_length) * 8, 256) + layout_for(total_length).total + 1024;
  return B200_OK;
}

int b200_sort_indices_chunked(b200_ctx* ctx, const b200_col* chunks, int n_chunks, int order, uint64_t* out_indices,
                              void* scratch, size_t scratch_bytes) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (n_chunks < 0 || (n_chunks > 0 && !chunks)) return b200_fail(ctx, B200_INVALID, "sort_indices_chunked: bad chunk list");
  int64_t total = 0;
  for (int c = 0; c < n_chunks; ++c) {
    if (chunks[c].length < 0) return b200_fail(ctx, B200_INVALID, "sort_indices_chunked: negative length");
    if (chunks[c].type != chunks[0].type) return b200_fail(ctx, B200_TYPE_ERROR, "sort_indices_chunked: chunk types differ");
    total += chunks[c].length;
  }
  if (total == 0) return B200_OK;
  size_t need = 0;
  b200_sort_indices_chunked_scratch_bytes(total, &need);
  if (!out_indices || !scratch || scratch_bytes < need) return b200_fail(ctx, B200_INVALID, "sort_indices_chunked: null output or scratch too small");
  char* base = reinterpret_cast<char*>(b200_align_up(reinterpret_cast<uintptr_t>(scratch), 256));
  const size_t arr = b200_align_up(size_t(total) * 8, 256);
  uint64_t* keys = reinterpret_cast<uint64_t*>(base);
  uint64_t* rows = reinterpret_cast<uint64_t*>(base + arr);
  uint64_t* skeys = reinterpret_cast<uint64_t*>(base + 2 * arr);
  uint64_t* srows = reinterpret_cast<uint64_t*>(base + 3 * arr);
  void* pair_scratch = base + 4 * arr;
  const size_t pair_bytes = scratch_bytes - (size_t(base - reinterpret_cast<char*>(scratch)) + 4 * arr);
  // the logical concatenation (ChunkedArraySorter's global indices, vector_sort.cc:657-832)
  int64_t off = 0;
  for (int c = 0; c < n_chunks; ++c) {
    if (chunks[c].length == 0) continue;
    B200_RETURN_NOT_OK(b200_sort_keys_rows(ctx, &chunks[c], order, uint64_t(off), keys + off, rows + off));
    off += chunks[c].length;
  }
  B200_RETURN_NOT_OK(b200_sort_pairs(ctx, keys, rows, total, skeys, srows, pair_scratch, pair_bytes));
  const unsigned grid = unsigned((total + 255) / 256);
  chunk_class_keys_kernel<<<grid, 256, 0, ctx->stream>>>(skeys, srows, total, keys);
  B200_LAUNCHED(ctx);
  B200_RETURN_NOT_OK(b200_sort_pairs(ctx, keys, srows, total, nullptr, out_indices, pair_scratch, pair_bytes));
  chunk_strip_kernel<<<grid, 256, 0, ctx->stream>>>(out_indices, total);
  B200_LAUNCHED(ctx);
  return B200_OK;
}

int b200_sort_indices_multi_scratch_bytes(int64_t length, size_t* out_bytes) {
  if (!out_bytes || length < 0) return B200_INVALID;
  // value keys (8 B) and payloads (4 B) of the key being sorted + the pair sort's own scratch
  *out_bytes = b200_align_up(size_t(length) * 8, 256) + b200_align_up(size_t(length) * 4, 256) + layout_for(length).total + 1024;
  return B200_OK;
}

int b200_sort_indices_multi(b200_ctx* ctx, const b200_col* keys, const int* orders, int n_keys, uint64_t* out_indices,
                            void* scratch, size_t scratch_bytes) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (n_keys < 1) return b200_fail(ctx, B200_INVALID, "Must specify one or more sort keys");   // vector_sort.cc:1636
  if (!keys || !orders) return b200_fail(ctx, B200_INVALID, "sort_indices_multi: null argument");
  if (n_keys == 1) return b200_sort_indices(ctx, &keys[0], orders[0], 0, out_indices, scratch, scratch_bytes);
  const int64_t n = keys[0].length;
  std::vector<SortParams> src(n_keys);
  for (int k = 0; k < n_keys; ++k) {
    memset(&src[k], 0, sizeof(SortParams));
    B200_RETURN_NOT_OK(fill_column_source(ctx, &keys[k], orders[k], &src[k]));
    if (keys[k].length != n) return b200_fail(ctx, B200_INVALID, "sort_indices_multi: sort keys of different lengths");
  }
  if (n == 0) return B200_OK;
  if (n > kMaxRows)
    return b200_fail(ctx, B200_NOT_IMPLEMENTED, "sort: %lld rows exceed the per-call limit of 2^30-1; shard the column", (long long)n);
  size_t need = 0;
  b200_sort_indices_multi_scratch_bytes(n, &need);
  if (!out_indices || !scratch || scratch_bytes < need) return b200_fail(ctx, B200_INVALID, "sort_indices_multi: null output or scratch too small");
  char* base = reinterpret_cast<char*>(b200_align_up(reinterpret_cast<uintptr_t>(scratch), 256));
  uint64_t* vkeys = reinterpret_cast<uint64_t*>(base);
  uint32_t* pay = reinterpret_cast<uint32_t*>(base + b200_align_up(size_t(n) * 8, 256));
  char* pair_scratch = base + b200_align_up(size_t(n) * 8, 256) + b200_align_up(size_t(n) * 4, 256);
  const size_t pair_bytes = scratch_bytes - size_t(pair_scratch - reinterpret_cast<char*>(scratch));
  const unsigned grid = unsigned((n + 255) / 256);
  // out_indices doubles as the permutation between keys: every kernel that reads it has finished (stream
  // order) before the sort that rewrites it starts
  // the least significant key sees the rows in row order: that is the single-column sort as it stands (keys
  // read straight from the column, NaN / null bins inside its class pass)
  B200_RETURN_NOT_OK(b200_sort_indices(ctx, &keys[n_keys - 1], orders[n_keys - 1], 0, out_indices, pair_scratch, pair_bytes));
  for (int k = n_keys - 2; k >= 0; --k) {
    mk_keys_kernel<<<grid, 256, 0, ctx->stream>>>(src[k], out_indices, vkeys, pay);
    B200_LAUNCHED(ctx);
    B200_RETURN_NOT_OK(b200_sort_pairs32(ctx, vkeys, pay, n, nullptr, out_indices, pair_scratch, pair_bytes));
    if (src[k].cls_digit >= 0) {   // the column can hold nulls or NaNs
      mk_class_kernel<<<grid, 256, 0, ctx->stream>>>(out_indices, n, vkeys, pay);
      B200_LAUNCHED(ctx);
      B200_RETURN_NOT_OK(b200_sort_pairs32(ctx, vkeys, pay, n, nullptr, out_indices, pair_scratch, pair_bytes));
    }
  }
  return B200_OK;
}

int b200_sort_indices_multi_chunked_scratch_bytes(int64_t total_length, size_t* out_bytes) {
  if (!out_bytes || total_length < 0) return B200_INVALID;
  size_t chunked = 0, multi = 0;
  b200_sort_indices_chunked_scratch_bytes(total_length, &chunked);
  b200_sort_indices_multi_scratch_bytes(total_length, &multi);
  // + the flat value keys (8 B) and classes (1 B) of the key being sorted
  multi += b200_align_up(size_t(total_length) * 8, 256) + b200_align_up(size_t(total_length), 256);
  *out_bytes = std::max(chunked, multi);
  return B200_OK;
}

int b200_sort_indices_multi_chunked(b200_ctx* ctx, const b200_col* keys, const int* orders, int n_keys, int n_chunks,
                                    uint64_t* out_indices, void* scratch, size_t scratch_bytes) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (n_keys < 1) return b200_fail(ctx, B200_INVALID, "Must specify one or more sort keys");
  if (n_chunks < 0 || !orders || (n_chunks > 0 && !keys)) return b200_fail(ctx, B200_INVALID, "sort_indices_multi_chunked: bad argument");
  if (n_chunks == 0) return B200_OK;
  if (n_chunks == 1) return b200_sort_indices_multi(ctx, keys, orders, n_keys, out_indices, scratch, scratch_bytes);
  if (n_keys == 1) return b200_sort_indices_chunked(ctx, keys, n_chunks, orders[0], out_indices, scratch, scratch_bytes);
  int64_t n = 0;
  for (int c = 0; c < n_chunks; ++c) n += keys[c].length;
  std::vector<SortParams> src(size_t(n_keys) * n_chunks);
  for (int k = 0; k < n_keys; ++k) {
    for (int c = 0; c < n_chunks; ++c) {
      const b200_col& col = keys[k * n_chunks + c];
      SortParams& p = src[size_t(k) * n_chunks + c];
      memset(&p, 0, sizeof(SortParams));
      if (col.type != keys[k * n_chunks].type) return b200_fail(ctx, B200_TYPE_ERROR, "sort_indices_multi_chunked: chunk types differ");
      B200_RETURN_NOT_OK(fill_column_source(ctx, &col, orders[k], &p));
      if (col.length != keys[c].length) return b200_fail(ctx, B200_INVALID, "sort_indices_multi_chunked: sort keys chunked differently");
    }
  }
  if (n == 0) return B200_OK;
  if (n > kMaxRows)
    return b200_fail(ctx, B200_NOT_IMPLEMENTED, "sort: %lld rows exceed the per-call limit of 2^30-1; shard the column", (long long)n);
  size_t need = 0;
  b200_sort_indices_multi_chunked_scratch_bytes(n, &need);
  if (!out_indices || !scratch || scratch_bytes < need) return b200_fail(ctx, B200_INVALID, "sort_indices_multi_chunked: null output or scratch too small");
  // least significant key: the chunked single-column sort (global indices over the concatenation)
  B200_RETURN_NOT_OK(b200_sort_indices_chunked(ctx, &keys[(n_keys - 1) * n_chunks], n_chunks, orders[n_keys - 1], out_indices, scratch, scratch_bytes));
  char* base = reinterpret_cast<char*>(b200_align_up(reinterpret_cast<uintptr_t>(scratch), 256));
  const size_t a8 = b200_align_up(size_t(n) * 8, 256), a4 = b200_align_up(size_t(n) * 4, 256), a1 = b200_align_up(size_t(n), 256);
  uint64_t* vkeys = reinterpret_cast<uint64_t*>(base);
  uint32_t* pay = reinterpret_cast<uint32_t*>(base + a8);
  uint64_t* flat_keys = reinterpret_cast<uint64_t*>(base + a8 + a4);
  uint8_t* flat_cls = reinterpret_cast<uint8_t*>(base + a8 + a4 + a8);
  char* pair_scratch = base + a8 + a4 + a8 + a1;
  const size_t pair_bytes = scratch_bytes - size_t(pair_scratch - reinterpret_cast<char*>(scratch));
  const unsigned grid = unsigned((n + 255) / 256);
  for (int k = n_keys - 2; k >= 0; --k) {
    int64_t off = 0;
    bool has_class = false;
    for (int c = 0; c < n_chunks; ++c) {
      const SortParams& p = src[size_t(k) * n_chunks + c];
      if (p.n == 0) continue;
      has_class = has_class || p.cls_digit >= 0;
      mk_flat_kernel<<<unsigned((p.n + 255) / 256), 256, 0, ctx->stream>>>(p, off, flat_keys, flat_cls);
      B200_LAUNCHED(ctx);
      off += p.n;
    }
    mk_gather_flat_kernel<<<grid, 256, 0, ctx->stream>>>(flat_keys, flat_cls, out_indices, n, vkeys, pay);
    B200_LAUNCHED(ctx);
    B200_RETURN_NOT_OK(b200_sort_pairs32(ctx, vkeys, pay, n, nullptr, out_indices, pair_scratch, pair_bytes));
    if (has_class) {
      mk_class_kernel<<<grid, 256, 0, ctx->stream>>>(out_indices, n, vkeys, pay);
      B200_LAUNCHED(ctx);
      B200_RETURN_NOT_OK(b200_sort_pairs32(ctx, vkeys, pay, n, nullptr, out_indices, pair_scratch, pair_bytes));
    }
  }
  return B200_OK;
}

int b200_sort_partition_scratch_bytes(int64_t length, size_t* out_bytes) {
  if (!out_bytes || length < 0) return B200_INVALID;
  const size_t n_tiles = size_t((length + kSpTile - 1) / kSpTile);
  *out_bytes = b200_align_up(n_tiles * kSpMax * sizeof(uint32_t), 256) + 2 * kSpMax * sizeof(int64_t) + 512;
  return B200_OK;
}

namespace {
struct PartitionPlan {
  SortParams p;
  Splitters sp;
  int64_t n_tiles;
  uint32_t* tile_counts;
  int64_t* totals;
  uint32_t* done;
};

int make_partition_plan(b200_ctx* ctx, const char* who, const b200_col* values, int order, const uint64_t* splitters,
                        int n_parts, void* scratch, size_t scratch_bytes, PartitionPlan* pl) {
  if (n_parts < 1 || n_parts > kSpMax || (n_parts > 1 && !splitters))
    return b200_fail(ctx, B200_INVALID, "%s: bad arguments (1..%d partitions)", who, kSpMax);
  memset(pl, 0, sizeof(*pl));
  B200_RETURN_NOT_OK(fill_column_source(ctx, values, order, &pl->p));
  if (pl->p.n >= (int64_t(1) << 32)) return b200_fail(ctx, B200_NOT_IMPLEMENTED, "%s: more than 2^32-1 rows per call", who);
  size_t need = 0;
  b200_sort_partition_scratch_bytes(pl->p.n, &need);
  if (pl->p.n > 0 && (!scratch || scratch_bytes < need)) return b200_fail(ctx, B200_INVALID, "%s: scratch too small", who);
  pl->sp.n_parts = n_parts;
  for (int j = 0; j + 1 < n_parts; ++j) {
    pl->sp.s[j] = splitters[j];
    if (j > 0 && pl->sp.s[j] < pl->sp.s[j - 1]) return b200_fail(ctx, B200_INVALID, "%s: splitters must ascend", who);
  }
  pl->n_tiles = (pl->p.n + kSpTile - 1) / kSpTile;
  char* base = reinterpret_cast<char*>(b200_align_up(reinterpret_cast<uintptr_t>(scratch), 256));
  pl->tile_counts = reinterpret_cast<uint32_t*>(base);
  pl->totals = reinterpret_cast<int64_t*>(base + b200_align_up(size_t(pl->n_tiles) * kSpMax * sizeof(uint32_t), 256));
  pl->done = reinterpret_cast<uint32_t*>(pl->totals + 2 * kSpMax);
  return B200_OK;
}

int partition_count(b200_ctx* ctx, const PartitionPlan& pl) {
  B200_TIMED(ctx, "sort_partition", (sp_count_kernel<<<(unsigned)pl.n_tiles, kSpThreads, 0, ctx->stream>>>(pl.p, pl.sp, pl.tile_counts)));
  B200_CUDA(ctx, cudaMemsetAsync(pl.done, 0, sizeof(uint32_t), ctx->stream));
  sp_scan_kernel<<<pl.sp.n_parts, 1024, 0, ctx->stream>>>(pl.tile_counts, pl.n_tiles, pl.sp.n_parts, pl.totals, pl.done);
  B200_LAUNCHED(ctx);
  return B200_OK;
}

int partition_counts_to_host(b200_ctx* ctx, const PartitionPlan& pl, int64_t* out_counts) {
  B200_CUDA(ctx, cudaMemcpyAsync(ctx->host_words + 32, pl.totals, kSpMax * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  B200_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  for (int q = 0; q < pl.sp.n_parts; ++q) out_counts[q] = ctx->host_words[32 + q];
  return B200_OK;
}
}  // namespace

int b200_sort_range_partition(b200_ctx* ctx, const b200_col* values, int order, uint64_t index_base,
                              const uint64_t* splitters, int n_parts, uint64_t* out_keys, uint64_t* out_rows,
                              int64_t* out_counts, void* scratch, size_t scratch_bytes) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (!out_counts) return b200_fail(ctx, B200_INVALID, "sort_range_partition: null out_counts");
  PartitionPlan pl;
  B200_RETURN_NOT_OK(make_partition_plan(ctx, "sort_range_partition", values, order, splitters, n_parts, scratch, scratch_bytes, &pl));
  for (int q = 0; q < n_parts; ++q) out_counts[q] = 0;
  if (pl.p.n == 0) return B200_OK;
  if (index_base + uint64_t(pl.p.n) > (uint64_t(1) << 62)) return b200_fail(ctx, B200_INDEX_ERROR, "sort_range_partition: row numbers beyond 2^62");
  if (!out_keys || !out_rows) return b200_fail(ctx, B200_INVALID, "sort_range_partition: null output");
  pl.p.index_base = index_base;
  B200_RETURN_NOT_OK(partition_count(ctx, pl));
  PushDst dst;
  memset(&dst, 0, sizeof(dst));
  for (int q = 0; q < n_parts; ++q) {
    dst.keys[q] = out_keys;
    dst.rows[q] = out_rows;
  }
  dst.totals = pl.totals;
  B200_TIMED(ctx, "sort_partition", (sp_scatter_kernel<<<(unsigned)pl.n_tiles, kSpThreads, 0, ctx->stream>>>(pl.p, pl.sp, pl.tile_counts, dst)));
  return partition_counts_to_host(ctx, pl, out_counts);
}

int b200_sort_range_partition_count(b200_ctx* ctx, const b200_col* values, int order, const uint64_t* splitters,
                                    int n_parts, int64_t* out_counts, void* scratch, size_t scratch_bytes) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (!out_counts) return b200_fail(ctx, B200_INVALID, "sort_range_partition_count: null out_counts");
  PartitionPlan pl;
  B200_RETURN_NOT_OK(make_partition_plan(ctx, "sort_range_partition_count", values, order, splitters, n_parts, scratch, scratch_bytes, &pl));
  for (int q = 0; q < n_parts; ++q) out_counts[q] = 0;
  if (pl.p.n == 0) return B200_OK;
  B200_RETURN_NOT_OK(partition_count(ctx, pl));
  return partition_counts_to_host(ctx, pl, out_counts);
}

int b200_sort_range_partition_push(b200_ctx* ctx, const b200_col* values, int order, uint64_t index_base,
                                   const uint64_t* splitters, int n_parts, uint64_t* const* dst_keys,
                                   void* const* dst_rows, const int64_t* dst_offsets, int rows32, void* scratch,
                                   size_t scratch_bytes) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (!dst_keys || !dst_rows || !dst_offsets) return b200_fail(ctx, B200_INVALID, "sort_range_partition_push: null destination table");
  PartitionPlan pl;
  B200_RETURN_NOT_OK(make_partition_plan(ctx, "sort_range_partition_push", values, order, splitters, n_parts, scratch, scratch_bytes, &pl));
  if (pl.p.n == 0) return B200_OK;
  const uint64_t row_limit = rows32 ? (uint64_t(1) << 30) : (uint64_t(1) << 62);
  if (index_base + uint64_t(pl.p.n) > row_limit)
    return b200_fail(ctx, B200_INDEX_ERROR, "sort_range_partition_push: row numbers beyond 2^%d", rows32 ? 30 : 62);
  pl.p.index_base = index_base;
  PushDst dst;
  memset(&dst, 0, sizeof(dst));
  for (int q = 0; q < n_parts; ++q) {
    if (!dst_keys[q] || !dst_rows[q] || dst_offsets[q] < 0) return b200_fail(ctx, B200_INVALID, "sort_range_partition_push: bad destination %d", q);
    dst.keys[q] = dst_keys[q];
    dst.rows[q] = dst_rows[q];
    dst.base[q] = dst_offsets[q];
  }
  dst.rows32 = rows32 ? 1 : 0;
  B200_TIMED(ctx, "sort_partition", (sp_scatter_kernel<<<(unsigned)pl.n_tiles, kSpThreads, 0, ctx->stream>>>(pl.p, pl.sp, pl.tile_counts, dst)));
  return B200_OK;
}

int b200_peer_alloc(b200_ctx* ctx, size_t bytes, void** out_ptr) {
  B200_ENTER(ctx);
  if (!ctx || !out_ptr) return B200_INVALID;
  *out_ptr = nullptr;
  B200_CUDA(ctx, cudaSetDevice(ctx->device));
  B200_CUDA(ctx, cudaMalloc(out_ptr, bytes ? bytes : 256));
  return B200_OK;
}
int b200_peer_free(b200_ctx* ctx, void* ptr) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (ptr) B200_CUDA(ctx, cudaFree(ptr));
  return B200_OK;
}
int b200_peer_export(b200_ctx* ctx, void* ptr, unsigned char* out_handle) {
  B200_ENTER(ctx);
  if (!ctx || !ptr || !out_handle) return B200_INVALID;
  static_assert(sizeof(cudaIpcMemHandle_t) == B200_PEER_HANDLE_BYTES, "IPC handle size");
  cudaIpcMemHandle_t h;
  B200_CUDA(ctx, cudaIpcGetMemHandle(&h, ptr));
  memcpy(out_handle, &h, sizeof(h));
  return B200_OK;
}
int b200_peer_open(b200_ctx* ctx, const unsigned char* handle, void** out_ptr) {
  B200_ENTER(ctx);
  if (!ctx || !handle || !out_ptr) return B200_INVALID;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  *out_ptr = nullptr;
  B200_CUDA(ctx, cudaSetDevice(ctx->device));
  B200_CUDA(ctx, cudaIpcOpenMemHandle(out_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return B200_OK;
}
int b200_peer_close(b200_ctx* ctx, void* ptr) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (ptr) B200_CUDA(ctx, cudaIpcCloseMemHandle(ptr));
  return B200_OK;
}

int b200_sort_keys(b200_ctx* ctx, const b200_col* values, int order, uint64_t* out_keys) {
  B200_ENTER(ctx);
  return b200_sort_keys_rows(ctx, values, order, 0, out_keys, nullptr);
}

int b200_sort_pairs(b200_ctx* ctx, const uint64_t* keys, const uint64_t* payload, int64_t length,
                    uint64_t* out_keys, uint64_t* out_payload, void* scratch, size_t scratch_bytes) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (length < 0) return b200_fail(ctx, B200_INVALID, "sort_pairs: negative length");
  if (length == 0) return B200_OK;
  if (!keys || !out_payload) return b200_fail(ctx, B200_INVALID, "sort_pairs: null argument");
  SortParams p;
  memset(&p, 0, sizeof(p));
  p.keys_in = keys;
  p.n = length;
  p.n_digits = kDigits;
  p.cls_digit = -1;
  p.wmask = ~uint64_t(0);
  p.gather = payload;  // NULL => emits the permutation itself
  p.out_idx = out_payload;
  p.out_keys = out_keys;
  return run_sort(ctx, p, scratch, scratch_bytes);
}

int b200_sort_pairs_bits(b200_ctx* ctx, const uint64_t* keys, const uint64_t* payload, int64_t length, int key_bits,
                         uint64_t* out_payload, void* scratch, size_t scratch_bytes) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (length < 0 || key_bits < 1 || key_bits > 64) return b200_fail(ctx, B200_INVALID, "sort_pairs_bits: bad argument");
  if (length == 0) return B200_OK;
  if (!keys || !out_payload) return b200_fail(ctx, B200_INVALID, "sort_pairs_bits: null argument");
  SortParams p;
  memset(&p, 0, sizeof(p));
  p.keys_in = keys;
  p.n = length;
  p.n_digits = (key_bits + 7) / 8;   // the digits above are zero by the caller's word: not read, not sorted
  p.cls_digit = -1;
  p.wmask = ~uint64_t(0);
  p.gather = payload;
  p.out_idx = out_payload;
  return run_sort(ctx, p, scratch, scratch_bytes);
}

int b200_sort_pairs32(b200_ctx* ctx, const uint64_t* keys, const uint32_t* payload32, int64_t length,
                      uint64_t* out_keys, uint64_t* out_payload, void* scratch, size_t scratch_bytes) {
  B200_ENTER(ctx);
  if (!ctx) return B200_INVALID;
  if (length < 0) return b200_fail(ctx, B200_INVALID, "sort_pairs32: negative length");
  if (length == 0) return B200_OK;
  if (!keys || !payload32 || !out_payload) return b200_fail(ctx, B200_INVALID, "sort_pairs32: null argument");
  SortParams p;
  memset(&p, 0, sizeof(p));
  p.keys_in = keys;
  p.pay_in = payload32;
  p.n = length;
  p.n_digits = kDigits;
  p.cls_digit = -1;
  p.wmask = ~uint64_t(0);
  p.out_idx = out_payload;
  p.out_keys = out_keys;
  return run_sort(ctx, p, scratch, scratch_bytes);
}

}  // extern "C"
