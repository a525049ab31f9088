// High-cardinality consume: radix-partition the rows by hash, then aggregate every partition in shared memory.
// (Included by groupby.cu inside its anonymous namespace, after the key / value loaders.)
//
// GrouperImpl::Consume + GroupedSumImpl / GroupedCountImpl::Consume (reference hash_aggregate.cc:358-408, 466-622)
// at a cardinality where the table no longer fits in L2: one row per thread with atomics on a multi-GB table costs
// 2-4 random 32-byte sector round trips per row (C4: 137 ms per 1 B rows, 1.8 % of the HBM peak).  Here the rows
// stream instead:
//   gb_hp_hist_kernel<1>   keys -> counts per level-1 partition (hash bits just below the table's index width, so
//                          a partition's groups later land in one contiguous region of the table)
//   gb_hp_part_kernel<1>   (key, value, row | null << 31) tuples scattered into the level-1 partitions: a 4096-row
//                          tile is ranked per bin with native shared 32-bit atomics, grouped in shared memory, and
//                          every bin's run leaves as contiguous stores at a position claimed from the bin's cursor
//                          (order inside a partition is immaterial: the earliest row is carried, not implied)
//   gb_hp_hist_kernel<2> / gb_hp_part_kernel<2>   the same once more inside every level-1 partition
//   gb_hp_agg_kernel       one CTA per final partition (~1.5 K keys): open-addressing table in shared memory
//                          {key, sum, valid count, null count, earliest row}; the groups leave as slot images
//   gb_import_kernel       (groupby.cu) folds the images into the global table, partition by partition, i.e. region
//                          by region of the table
// Rows that cannot be placed (special keys, a shared table that fills up, no room for a partition's images) are
// flagged in the deferred-row bitmap and taken by the generic kernel, like in the other fast paths.
// HBM bytes per row at C4 (int64 key, float64 value): 8 + (16 + 20) + 8 + (20 + 20) + 20 = 132, + the images and one
// sweep over the table.

// tile shape of the partition kernels (a tuning sweep builds variants with -D)
#ifndef B200_HP_ITEMS
#define B200_HP_ITEMS 8
#endif
#ifndef B200_HP_PART_MINBLOCKS
#define B200_HP_PART_MINBLOCKS 2
#endif
constexpr int kHpThreads = 512;
constexpr int kHpItems = B200_HP_ITEMS;
constexpr int kHpTile = kHpThreads * kHpItems;   // 4096 rows
constexpr int kHpMaxBins = 512;
constexpr int kHpSlots = 4096;                   // shared-memory table of the aggregation kernel
constexpr int kHpKeysPerPart = 1536;             // target distinct keys per final partition (load 3/8)
constexpr int kHpProbe = 48;
constexpr uint32_t kHpNullBit = 0x80000000u;
constexpr uint32_t kHpRowMask = 0x7FFFFFFFu;

struct HpBufs {
  uint64_t* keys;
  uint64_t* vals;
  uint32_t* rows;   // row inside the batch | value-is-null << 31
  uint32_t* ord;    // nullable: explicit global row number - the batch's smallest one (b200_groupby_consume_rows);
                    // the earliest row of a group is then the minimum of these instead of the minimum of `rows`
};

struct HpArgs {
  HpBufs src;             // level 2 input (level 1 reads the columns)
  HpBufs dst;
  const uint32_t* base;   // level 2: [parts + 1] starts of the level-1 partitions inside src
  int32_t parts;          // level 2: number of level-1 partitions (1 at level 1)
  uint32_t* hist;         // [parts * bins]
  uint32_t* cursor;       // [parts * bins] next free position in dst
  int32_t shift;          // bin = (mix64(key) >> shift) & (bins - 1)
  int32_t bins;
  int32_t has_val;
  uint64_t ord_base;      // level 1 with explicit row numbers: the smallest one
  int32_t fast;           // level 1: one key column of `fast` bytes (4 / 8) without nulls, 64-bit value column without
                          // nulls (or none): straight-line loads, all of a thread's rows in flight together
};

__device__ __forceinline__ bool hp_row_live(const GbArgs& a, int64_t row) {
  return !(a.select && !((a.select[row >> 5] >> (row & 31)) & 1u));
}

// partition whose range holds `pos` (base ascends; empty partitions are stepped over by the callers)
__device__ __forceinline__ uint32_t hp_find_part(const uint32_t* base, uint32_t parts, uint32_t pos) {
  uint32_t lo = 0, hi = parts;   // invariant: base[lo] <= pos < base[hi]
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (base[mid] <= pos) lo = mid;
    else hi = mid;
  }
  return lo;
}

// LEVEL 1 reads the key columns, LEVEL 2 the level-1 output.  Both walk a flat list of 4096-row tiles; at level 2
// a tile that straddles level-1 partitions is taken segment by segment, so a skewed partition is still spread over
// the whole grid.
template <int LEVEL>
__global__ void __launch_bounds__(kHpThreads) gb_hp_hist_kernel(const __grid_constant__ GbArgs a, const __grid_constant__ HpArgs h) {
  __shared__ uint32_t cnt[kHpMaxBins];
  __shared__ uint32_t sbase[kHpMaxBins + 1];   // level 2: the partition starts (searched once per tile by every thread)
  const int t = threadIdx.x;
  const uint32_t mask = uint32_t(h.bins - 1);
  if (LEVEL == 2) {
    for (int i = t; i <= h.parts; i += kHpThreads) sbase[i] = h.base[i];
    __syncthreads();
  }
  const int64_t n = LEVEL == 2 ? int64_t(sbase[h.parts]) : a.n;
  uint32_t held = 0xFFFFFFFFu;   // partition whose counts sit in shared memory
  for (int64_t tile0 = int64_t(blockIdx.x) * kHpTile; tile0 < n; tile0 += int64_t(gridDim.x) * kHpTile) {
    const int64_t tile1 = min(n, tile0 + kHpTile);
    int64_t seg0 = tile0;
    uint32_t p = LEVEL == 2 ? hp_find_part(sbase, uint32_t(h.parts), uint32_t(tile0)) : 0;
    while (seg0 < tile1) {
      const int64_t seg1 = LEVEL == 2 ? min(tile1, int64_t(sbase[p + 1])) : tile1;
      if (seg1 > seg0) {
        if (p != held) {
          __syncthreads();
          for (int i = t; i < h.bins; i += kHpThreads) {
            if (held != 0xFFFFFFFFu) {
              const uint32_t c = cnt[i];
              if (c) atomicAdd(h.hist + size_t(held) * h.bins + i, c);
            }
            cnt[i] = 0;
          }
          held = p;
          __syncthreads();
        }
        if (LEVEL == 2 || h.fast) {
          uint64_t key[kHpItems];
          const uint64_t* k64 = LEVEL == 2 ? h.src.keys : reinterpret_cast<const uint64_t*>(a.keys[0].data);
          const uint32_t* k32 = reinterpret_cast<const uint32_t*>(a.keys[0].data);
          if (LEVEL == 2 || h.fast == 8) {
#pragma unroll
            for (int i = 0; i < kHpItems; ++i) {
              const int64_t row = seg0 + i * kHpThreads + t;
              key[i] = row < seg1 ? k64[row] : 0;
            }
          } else {
#pragma unroll
            for (int i = 0; i < kHpItems; ++i) {
              const int64_t row = seg0 + i * kHpThreads + t;
              key[i] = row < seg1 ? uint64_t(k32[row]) : 0;
            }
          }
#pragma unroll
          for (int i = 0; i < kHpItems; ++i) {
            const int64_t row = seg0 + i * kHpThreads + t;
            bool live = row < seg1;
            if (LEVEL == 1 && live) live = hp_row_live(a, row) && !(h.fast == 8 && key[i] == kEmpty);
            if (live) atomicAdd(&cnt[uint32_t(mix64(key[i]) >> h.shift) & mask], 1u);
          }
        } else {
#pragma unroll
          for (int i = 0; i < kHpItems; ++i) {
            const int64_t row = seg0 + i * kHpThreads + t;
            if (row < seg1 && hp_row_live(a, row)) {
              int special;
              const uint64_t key = load_key(a, row, &special);
              if (!special) atomicAdd(&cnt[uint32_t(mix64(key) >> h.shift) & mask], 1u);
            }
          }
        }
      }
      seg0 = seg1;
      ++p;
    }
  }
  __syncthreads();
  if (held != 0xFFFFFFFFu) {
    for (int i = t; i < h.bins; i += kHpThreads) {
      const uint32_t c = cnt[i];
      if (c) atomicAdd(h.hist + size_t(held) * h.bins + i, c);
    }
  }
}

// exclusive scan of `count` words (one CTA): starts -> cursor[i] and base[i]; base[count] = total
__global__ void __launch_bounds__(1024) gb_hp_scan_kernel(const uint32_t* hist, uint32_t count, uint32_t* cursor, uint32_t* base) {
  __shared__ uint32_t warp_sums[32];
  __shared__ uint32_t carry_s;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  if (t == 0) carry_s = 0;
  __syncthreads();
  for (uint32_t b0 = 0; b0 < count; b0 += 1024) {
    const uint32_t i = b0 + t;
    const uint32_t c = i < count ? hist[i] : 0;
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
    if (i < count) {
      if (cursor) cursor[i] = carry + wpre + incl - c;
      base[i] = carry + wpre + incl - c;
    }
    __syncthreads();
    if (t == 1023) carry_s = carry + wpre + incl;
    __syncthreads();
  }
  if (t == 0) base[count] = carry_s;
}

struct HpPartSmem {
  uint64_t keys[kHpTile];
  uint64_t vals[kHpTile];
  uint32_t rows[kHpTile];
  uint16_t bin[kHpTile];
  uint32_t cnt[kHpMaxBins];
  uint32_t off[kHpMaxBins];
  uint32_t gbase[kHpMaxBins];
  uint32_t warp_sums[kHpThreads / 32];
  uint32_t sbase[kHpMaxBins + 1];
  uint32_t deferred;
  uint32_t ord[kHpTile];   // last: only the kernels that carry explicit row numbers allocate it
};

template <int LEVEL, bool kOrd>
__global__ void __launch_bounds__(kHpThreads, B200_HP_PART_MINBLOCKS) gb_hp_part_kernel(const __grid_constant__ GbArgs a, const __grid_constant__ HpArgs h) {
  extern __shared__ __align__(16) unsigned char hp_smem_raw[];
  HpPartSmem& S = *reinterpret_cast<HpPartSmem*>(hp_smem_raw);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const uint32_t mask = uint32_t(h.bins - 1);
  if (LEVEL == 2) {
    for (int i = t; i <= h.parts; i += kHpThreads) S.sbase[i] = h.base[i];
    __syncthreads();
  }
  const int64_t n = LEVEL == 2 ? int64_t(S.sbase[h.parts]) : a.n;
  if (t == 0) S.deferred = 0;
  for (int64_t tile0 = int64_t(blockIdx.x) * kHpTile; tile0 < n; tile0 += int64_t(gridDim.x) * kHpTile) {
    const int64_t tile1 = min(n, tile0 + kHpTile);
    int64_t seg0 = tile0;
    uint32_t p = LEVEL == 2 ? hp_find_part(S.sbase, uint32_t(h.parts), uint32_t(tile0)) : 0;
    while (seg0 < tile1) {
      const int64_t seg1 = LEVEL == 2 ? min(tile1, int64_t(S.sbase[p + 1])) : tile1;
      if (seg1 > seg0) {
        uint32_t* cursor = h.cursor + size_t(p) * h.bins;
        __syncthreads();   // the previous segment's stores have read S
        for (int i = t; i < h.bins; i += kHpThreads) S.cnt[i] = 0;
        __syncthreads();
        // ---- load, bin, rank inside the tile's bin (native ATOMS.ADD with return)
        uint64_t key[kHpItems], val[kHpItems];
        uint32_t r32[kHpItems], meta[kHpItems];   // meta = bin | rank << 16, 0xFFFFFFFF = not staged
        uint32_t ord[kOrd ? kHpItems : 1];
        constexpr bool has_ord = kOrd;
        if (kOrd) {
#pragma unroll
          for (int i = 0; i < kHpItems; ++i) {
            const int64_t row = seg0 + i * kHpThreads + t;
            ord[kOrd ? i : 0] = 0;
            if (row < seg1) ord[kOrd ? i : 0] = LEVEL == 2 ? h.src.ord[row] : uint32_t(a.row_ids[row] - h.ord_base);
          }
        }
        if (LEVEL == 2 || h.fast) {
          // straight-line loads: all of the thread's rows are requested before the first is looked at
          const uint64_t* k64 = LEVEL == 2 ? h.src.keys : reinterpret_cast<const uint64_t*>(a.keys[0].data);
          const uint32_t* k32 = reinterpret_cast<const uint32_t*>(a.keys[0].data);
          const uint64_t* vp = LEVEL == 2 ? h.src.vals : reinterpret_cast<const uint64_t*>(a.cols[0].data);
          if (LEVEL == 2 || h.fast == 8) {
#pragma unroll
            for (int i = 0; i < kHpItems; ++i) {
              const int64_t row = seg0 + i * kHpThreads + t;
              key[i] = row < seg1 ? k64[row] : 0;
            }
          } else {
#pragma unroll
            for (int i = 0; i < kHpItems; ++i) {
              const int64_t row = seg0 + i * kHpThreads + t;
              key[i] = row < seg1 ? uint64_t(k32[row]) : 0;
            }
          }
#pragma unroll
          for (int i = 0; i < kHpItems; ++i) {
            const int64_t row = seg0 + i * kHpThreads + t;
            val[i] = (h.has_val && row < seg1) ? vp[row] : 0;
            r32[i] = LEVEL == 2 ? (row < seg1 ? h.src.rows[row] : 0u) : uint32_t(row);
          }
#pragma unroll
          for (int i = 0; i < kHpItems; ++i) {
            const int64_t row = seg0 + i * kHpThreads + t;
            meta[i] = 0xFFFFFFFFu;
            bool live = row < seg1;
            if (LEVEL == 1 && live) {
              live = hp_row_live(a, row);
              if (live && h.fast == 8 && key[i] == kEmpty) {   // the key equal to the EMPTY sentinel: generic kernel
                atomicOr(a.defer_out + (row >> 5), 1u << (row & 31));
                atomicAdd(&S.deferred, 1u);
                live = false;
              }
            }
            if (live) meta[i] = uint32_t(mix64(key[i]) >> h.shift) & mask;
          }
        } else {
#pragma unroll
          for (int i = 0; i < kHpItems; ++i) {
            const int64_t row = seg0 + i * kHpThreads + t;
            meta[i] = 0xFFFFFFFFu;
            key[i] = 0;
            val[i] = 0;
            r32[i] = 0;
            if (row < seg1 && hp_row_live(a, row)) {
              int special;
              key[i] = load_key(a, row, &special);
              if (special) {   // the null key / the key equal to the EMPTY sentinel: left to the generic kernel
                atomicOr(a.defer_out + (row >> 5), 1u << (row & 31));
                atomicAdd(&S.deferred, 1u);
              } else {
                r32[i] = uint32_t(row);
                if (h.has_val) {
                  const Val v = load_val(a.cols[0], row);
                  val[i] = v.bits;
                  if (!v.valid) r32[i] |= kHpNullBit;
                }
                meta[i] = uint32_t(mix64(key[i]) >> h.shift) & mask;
              }
            }
          }
        }
#pragma unroll
        for (int i = 0; i < kHpItems; ++i)
          if (meta[i] != 0xFFFFFFFFu) meta[i] |= atomicAdd(&S.cnt[meta[i]], 1u) << 16;
        __syncthreads();
        // ---- claim every bin's run in the destination; exclusive scan of the tile's counts
        {
          const uint32_t c = t < h.bins ? S.cnt[t] : 0;
          if (t < h.bins) S.gbase[t] = c ? atomicAdd(cursor + t, c) : 0;
          uint32_t incl = c;
#pragma unroll
          for (int d = 1; d < 32; d <<= 1) {
            const uint32_t o = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += o;
          }
          if (lane == 31) S.warp_sums[warp] = incl;
          __syncthreads();
          uint32_t wpre = 0;
#pragma unroll
          for (int w = 0; w < kHpThreads / 32; ++w)
            if (w < warp) wpre += S.warp_sums[w];
          if (t < h.bins) S.off[t] = wpre + incl - c;
        }
        __syncthreads();
        const uint32_t staged = S.off[h.bins - 1] + S.cnt[h.bins - 1];
        // ---- group the tile by bin in shared memory
#pragma unroll
        for (int i = 0; i < kHpItems; ++i) {
          if (meta[i] != 0xFFFFFFFFu) {
            const uint32_t b = meta[i] & 0xFFFFu;
            const uint32_t pos = S.off[b] + (meta[i] >> 16);
            S.keys[pos] = key[i];
            S.vals[pos] = val[i];
            S.rows[pos] = r32[i];
            if (kOrd) S.ord[pos] = ord[kOrd ? i : 0];
            S.bin[pos] = uint16_t(b);
          }
        }
        __syncthreads();
        // ---- bin-contiguous stores
        for (uint32_t j = t; j < staged; j += kHpThreads) {
          const uint32_t b = S.bin[j];
          const uint32_t g = S.gbase[b] + (j - S.off[b]);
          h.dst.keys[g] = S.keys[j];
          if (h.has_val) h.dst.vals[g] = S.vals[j];
          h.dst.rows[g] = S.rows[j];
          if (has_ord) h.dst.ord[g] = S.ord[j];
        }
      }
      seg0 = seg1;
      ++p;
    }
  }
  if (LEVEL == 1) {
    __syncthreads();
    if (t == 0 && S.deferred) atomicAdd(a.counters + 1, S.deferred);
  }
}

struct HpAggSmem {
  uint64_t key[kHpSlots];
  uint64_t sum[kHpSlots];
  uint32_t cnt_valid[kHpSlots];
  uint32_t cnt_null[kHpSlots];
  uint32_t row[kHpSlots];
  uint8_t list[kHpThreads / 32][32];   // per warp: the lanes whose slot is in use, compacted
  uint8_t code[64];                    // per 64-bit word of a slot image: where its value comes from (kImg*)
  uint32_t part, begin, end;
  uint32_t image_base;
  uint32_t deferred;
  uint32_t n_keys;                     // slots claimed for the current work item
  uint32_t cursor;                     // images handed out so far (of n_keys)
};
enum { kImgZero = 0, kImgKey, kImgRow, kImgValid, kImgNull, kImgSum };
static_assert(2 * (sizeof(HpAggSmem) + 1024) <= 228 * 1024, "two aggregation CTAs per SM");

// sum += v on a 64-bit integer accumulator in shared memory with native 32-bit atomics: the carry out of the low
// word is known to the thread whose add produced it
__device__ __forceinline__ void hp_add_u64(uint64_t* acc, uint64_t v) {
  uint32_t* w = reinterpret_cast<uint32_t*>(acc);
  const uint32_t lo = uint32_t(v);
  const uint32_t old = atomicAdd(w, lo);
  const uint32_t hi = uint32_t(v >> 32) + (old + lo < old ? 1u : 0u);
  if (hi) atomicAdd(w + 1, hi);
}

// work items of the aggregation kernel: a partition of more than seg_rows rows (a hot key) is cut into segments that
// different CTAs aggregate; the import folds their images together
__global__ void gb_hp_items_kernel(const uint32_t* base, uint32_t n_parts, uint32_t seg_rows, uint32_t* items) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  // (at least one item per partition: while no partition is cut, item i IS partition i and needs no search)
  if (i < n_parts) items[i] = max(1u, (base[i + 1] - base[i] + seg_rows - 1) / seg_rows);
}

// kFloatSum: the value column sums as float64 (ATOMS.CAS.64 loop), else as a wrapping 64-bit integer
template <bool kFloatSum>
__global__ void __launch_bounds__(kHpThreads, 2) gb_hp_agg_kernel(const __grid_constant__ GbArgs a, const HpBufs src, const uint32_t* base,
                                                                  const uint32_t* item_base, const uint32_t n_parts,
                                                                  const uint32_t seg_rows, char* images, const uint32_t image_cap,
                                                                  uint32_t* ticket) {
  extern __shared__ __align__(16) unsigned char hp_smem_raw[];
  HpAggSmem& S = *reinterpret_cast<HpAggSmem*>(hp_smem_raw);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int words = a.s.stride / 8;
  if (t == 0) S.deferred = 0;
  if (t == 0 && blockIdx.x == 0) a.counters[7] = item_base[n_parts];   // the host wants to know if a partition was cut
  if (t < 64) {
    // layout of a slot image (gb_import_kernel / gb_emit_kernel read it): [key][first row][aggregate states]
    int c = t == 0 ? kImgKey : (t == 1 ? kImgRow : kImgZero);
    for (int j = 0; j < a.s.n_aggs; ++j) {
      const GbAgg& g = a.s.aggs[j];
      if (g.kind == B200_HASH_COUNT) {
        if (g.off / 8 == t) c = g.option ? kImgNull : kImgValid;
      } else {   // SUM / MEAN: {sum, count of valid values}
        if (g.off / 8 == t) c = kImgSum;
        if (g.off / 8 + 1 == t) c = kImgValid;
      }
    }
    S.code[t] = uint8_t(c);
  }
  for (;;) {
    __syncthreads();
    if (t == 0) {
      // the next work item: one thread walks the dependent loads while the others clear the table
      const uint32_t item = atomicAdd(ticket, 1u);
      const uint32_t n_items = item_base[n_parts];
      uint32_t b = 0, e = 0;
      if (item < n_items) {
        const uint32_t p = n_items == n_parts ? item : hp_find_part(item_base, n_parts, item);
        b = base[p] + (item - item_base[p]) * seg_rows;
        e = min(base[p + 1], b + seg_rows);
      }
      S.part = item < n_items ? item : 0xFFFFFFFFu;
      S.begin = b;
      S.end = e;
      S.n_keys = 0;
      S.cursor = 0;
    }
    for (int i = t; i < kHpSlots; i += kHpThreads) {
      S.key[i] = kEmpty;
      S.sum[i] = 0;
      S.cnt_valid[i] = 0;
      S.cnt_null[i] = 0;
      S.row[i] = 0xFFFFFFFFu;
    }
    __syncthreads();
    if (S.part == 0xFFFFFFFFu) break;
    const uint32_t begin = S.begin, end = S.end;
    constexpr int kU = 4;   // rows per thread in flight: their loads are issued together
    for (uint32_t i0 = begin; i0 < end; i0 += kHpThreads * kU) {
      uint64_t key[kU], v[kU];
      uint32_t r32[kU], ord[kU];
      bool in[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const uint32_t i = i0 + u * kHpThreads + t;
        in[u] = i < end;
        key[u] = in[u] ? src.keys[i] : 0;
        r32[u] = in[u] ? src.rows[i] : 0;
        ord[u] = (in[u] && src.ord) ? src.ord[i] : (r32[u] & kHpRowMask);
        v[u] = (in[u] && a.n_cols) ? src.vals[i] : 0;
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        __syncwarp();
        if (!in[u]) continue;
        uint32_t slot = uint32_t(mix64(key[u]) >> 50) & uint32_t(kHpSlots - 1);
        bool found = false;
        // (a warp's row costs the LONGEST probe sequence among its 32 lanes: the step is kept short — the hit first,
        // the claim only on an empty slot — and rolled, so that the address is one add and one mask)
#pragma unroll 1
        for (int probe = 0; probe < kHpProbe; ++probe) {
          uint64_t cur = *reinterpret_cast<volatile uint64_t*>(&S.key[slot]);
          if (cur == key[u]) {
            found = true;
            break;
          }
          if (cur == kEmpty) {
            cur = atomicCAS(reinterpret_cast<unsigned long long*>(&S.key[slot]), (unsigned long long)kEmpty, (unsigned long long)key[u]);
            if (cur == kEmpty) atomicAdd(&S.n_keys, 1u);
            if (cur == kEmpty || cur == key[u]) {
              found = true;
              break;
            }
          }
          slot = (slot + 1) & uint32_t(kHpSlots - 1);
        }
        const uint32_t row = r32[u] & kHpRowMask;
        if (!found) {   // more distinct keys than the shared table takes: the generic kernel aggregates this row
          atomicOr(a.defer_out + (row >> 5), 1u << (row & 31));
          atomicAdd(&S.deferred, 1u);
          continue;
        }
        atomicMin(&S.row[slot], ord[u]);
        if (r32[u] & kHpNullBit) {
          atomicAdd(&S.cnt_null[slot], 1u);
        } else {
          atomicAdd(&S.cnt_valid[slot], 1u);
          if (a.n_cols) {
            // (one row after the other on purpose: with the four rows' compare-and-swaps issued together — 2048 rows
            // of a CTA in flight over ~1500 groups — the window between reading a sum and swapping it grows, most
            // swaps lose and retry, and the kernel took 22.9 ms instead of 11.0 per 1 B rows)
            if (kFloatSum) atomicAdd(reinterpret_cast<double*>(&S.sum[slot]), __longlong_as_double(v[u]));
            else hp_add_u64(&S.sum[slot], v[u]);
          }
        }
      }
    }
    __syncthreads();
    // ---- the partition's groups leave as slot images (in any order: finalize ranks them by first row)
    if (t == 0) {
      // one atomicAdd reserves the whole item's images.  When the list is full the part of the claim that lies
      // inside it is filled with EMPTY keys, which the import skips.
      const uint32_t total = S.n_keys;
      uint32_t got = 0;
      if (total) {
        got = atomicAdd(a.counters + 6, total);
        if (uint64_t(got) + total > image_cap) {
          for (uint32_t i = got; i < image_cap; ++i) *reinterpret_cast<uint64_t*>(images + size_t(i) * a.s.stride) = kEmpty;
          got = 0xFFFFFFFFu;
        }
      }
      S.image_base = got;
    }
    __syncthreads();
    const uint32_t ibase = S.image_base;
    if (ibase == 0xFFFFFFFFu) {
      // no room for this partition's images: its rows go to the generic kernel
      for (uint32_t i = begin + t; i < end; i += kHpThreads) {
        const uint32_t row = src.rows[i] & kHpRowMask;
        atomicOr(a.defer_out + (row >> 5), 1u << (row & 31));
      }
      if (t == 0) S.deferred += end - begin;
      continue;
    }
    // A warp looks at 32 consecutive slots at a time, compacts the used ones into a list and writes their images
    // WORD by word across its lanes: consecutive lanes store consecutive 8-byte words (one thread per image and
    // a run-time walk over the aggregates took 28 % of the kernel, profiles/r02_gb_hp_agg_ncu_summary.txt).
    for (int grp = warp; grp < kHpSlots / 32; grp += kHpThreads / 32) {
      const int s = grp * 32 + lane;
      const bool used = S.key[s] != kEmpty;
      const unsigned m = __ballot_sync(0xffffffffu, used);
      if (m == 0) continue;
      const int cnt = __popc(m);
      uint32_t first = 0;
      if (lane == 0) first = atomicAdd(&S.cursor, uint32_t(cnt));
      first = __shfl_sync(0xffffffffu, first, 0);
      if (used) S.list[warp][__popc(m & ((1u << lane) - 1u))] = uint8_t(lane);
      __syncwarp();
      uint64_t* out = reinterpret_cast<uint64_t*>(images + size_t(ibase + first) * a.s.stride);
      const int n_words = cnt * words;
      for (int idx = lane; idx < n_words; idx += 32) {
        const int mi = idx / words, w = idx - mi * words;
        const int slot = grp * 32 + S.list[warp][mi];
        uint64_t v = 0;
        switch (S.code[w]) {
          case kImgKey: v = S.key[slot]; break;
          case kImgRow: v = a.row_base + uint64_t(S.row[slot]); break;   // with explicit row numbers row_base is the batch's smallest one
          case kImgValid: v = S.cnt_valid[slot]; break;
          case kImgNull: v = S.cnt_null[slot]; break;
          case kImgSum: v = S.sum[slot]; break;
          default: break;
        }
        out[idx] = v;
      }
      __syncwarp();
    }
  }
  if (t == 0 && S.deferred) atomicAdd(a.counters + 1, S.deferred);
}

// slot images -> global table (gb_import_kernel with the new-group count folded per CTA: at 10^8 groups one global
// counter bumped per group would serialise in L2)
__global__ void __launch_bounds__(256) gb_hp_import_kernel(const __grid_constant__ GbSchema s, const char* src, uint32_t n, GbTable dst,
                                                           uint32_t* counters) {
  __shared__ uint32_t s_new;
  if (threadIdx.x == 0) s_new = 0;
  __syncthreads();
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const char* ss = src + uint64_t(i) * s.stride;
    bool claimed = false;
    uint64_t slot_index;
    const uint64_t key = *reinterpret_cast<const uint64_t*>(ss);
    char* ds = key == kEmpty ? nullptr : global_slot(dst, s.stride, key, 0, true, &claimed, &slot_index);
    if (key == kEmpty) {
      // an unused entry of a fixed-size image list (gb_fused8_kernel)
    } else if (!ds) {
      atomicExch(counters + 3, 4u);
    } else {
      if (claimed) atomicAdd(&s_new, 1u);
      merge_state(ds, ss, s);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0 && s_new) atomicAdd(counters + 0, s_new);
}

// smallest and largest explicit row number of a batch -> mm[0], mm[1] (mm[0] starts at all ones, mm[1] at zero)
__global__ void gb_hp_minmax_kernel(const uint64_t* row_ids, int64_t n, unsigned long long* mm) {
  uint64_t lo = ~uint64_t(0), hi = 0;
  for (int64_t i = int64_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
    const uint64_t r = row_ids[i];
    lo = min(lo, r);
    hi = max(hi, r);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, d));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, d));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(mm, (unsigned long long)lo);
    atomicMax(mm + 1, (unsigned long long)hi);
  }
}
