// Fused compare -> filter -> derive -> aggregate in ONE pass over the inputs.
// (Included by groupby.cu inside its anonymous namespace, after the tiny kernels.)
//
// Upstream the Q1 shape runs as less_equal (scalar_compare.cc) -> Filter of every column
// (vector_selection.cc:1776-1866) -> multiply / subtract (scalar_arithmetic.cc:69-190) -> GroupBy
// (hash_aggregate.cc:890-966): every stage reads and writes whole columns.  gb_fused8_kernel reads the predicate
// column, the byte-wide key columns and every distinct float64 argument column once (Q1: 4 + 2 + 3 x 8 = 30 B/row),
// evaluates the predicate and the derived arguments `a op1 (scalar op2 b)` in registers with the same single IEEE
// operations b200_arithmetic uses (__dadd_rn / __dsub_rn / __dmul_rn: never contracted into an FMA), and keeps
// the <= 8 groups' sums (one per argument stream) and counts THREAD-PRIVATE in shared memory, laid out
// [group][stream][thread] so that a warp's 32 accumulators of one (group, stream) are consecutive words: plain
// LDS / DADD / STS, no atomics and no bank conflict whatever groups the lanes hold.  (The first version kept them in
// registers with predicated adds like gb_tiny8_kernel: 8 groups x 4 streams = 64 registers of accumulators, 128
// registers per thread, 16 warps per SM — 8.2 ms per 600 M rows, latency-bound at 0.34 of the HBM peak.)
// A CTA leaves its groups as slot images in a scratch list; nothing touches the table until the host has seen that
// no CTA met a 9th group — if one did, the call is redone through the unfused operations.

constexpr int kFusedStreams = 4;   // distinct aggregate arguments
constexpr int kFusedCols = 4;      // distinct float64 columns they read

struct FusedStream {
  int32_t kind;      // 0: column ia; 1: col[ia] op1 (scalar op2 col[ib])
  int32_t ia, ib;
  int32_t op1, op2;  // b200_arith_op
  int32_t pad;
  double scalar;
};

struct FusedArgs {
  const double* cols[kFusedCols];
  FusedStream streams[kFusedStreams];
  int32_t n_cols, n_streams;
  int32_t n_plain;         // streams [0, n_plain) are the columns [0, n_plain) as they are (no operand selection per
                           // row); the derived streams follow and pick their operands among all columns
  const void* pred_data;   // nullable: 4-byte predicate column (shifted by its offset)
  int32_t pred_type;       // B200_INT32 / B200_UINT32
  int32_t pred_op;         // kCmp* order of b200_compare_scalar: 0 eq, 1 ne, 2 gt, 3 ge, 4 lt, 5 le
  int64_t pred_scalar;
  char* images;            // [gridDim.x * kTinyGroups] slot images, key == EMPTY: unused entry
  uint32_t* overflow;      // set when a CTA met more than kTinyGroups keys
};

__device__ __forceinline__ double fused_arith(int op, double x, double y) {
  return op == B200_ADD ? __dadd_rn(x, y) : (op == B200_SUBTRACT ? __dsub_rn(x, y) : __dmul_rn(x, y));
}

__device__ __forceinline__ double fused_pick(int i, const double (&c)[kFusedCols]) {
  return i == 0 ? c[0] : (i == 1 ? c[1] : (i == 2 ? c[2] : c[3]));
}

template <typename T>
__device__ __forceinline__ bool fused_cmp(int op, T x, T s) {
  switch (op) {
    case 0: return x == s;
    case 1: return x != s;
    case 2: return x > s;
    case 3: return x >= s;
    case 4: return x < s;
    default: return x <= s;
  }
}

template <int NS>
struct FusedSmem {
  double acc[kTinyGroups][NS][kTinyThreads];
  uint32_t cnt[kTinyGroups][kTinyThreads];
  uint64_t keys[kTinyGroups];
  uint32_t first[kTinyGroups];
  double totals[kTinyGroups][NS];
  unsigned long long counts[kTinyGroups];
};

template <int NS>
__global__ void __launch_bounds__(kTinyThreads, 3) gb_fused8_kernel(const __grid_constant__ GbArgs a, const __grid_constant__ FusedArgs f) {
  extern __shared__ __align__(16) unsigned char smem[];
  FusedSmem<NS>& S = *reinterpret_cast<FusedSmem<NS>*>(smem);
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
#pragma unroll
  for (int q = 0; q < kTinyGroups; ++q) {
    S.cnt[q][t] = 0;
#pragma unroll
    for (int s = 0; s < NS; ++s) S.acc[q][s][t] = 0.0;
  }
  if (t < kTinyGroups) {
    S.keys[t] = kEmpty;
    S.first[t] = 0xFFFFFFFFu;
  }
  __syncthreads();
  const int64_t begin = int64_t(blockIdx.x) * a.rows_per_block;
  const int64_t end = min(a.n, begin + a.rows_per_block);
  const uint8_t* k0 = reinterpret_cast<const uint8_t*>(a.keys[0].data);
  const uint8_t* k1 = a.s.n_keys > 1 ? reinterpret_cast<const uint8_t*>(a.keys[1].data) : nullptr;
  const uint8_t* k2 = a.s.n_keys > 2 ? reinterpret_cast<const uint8_t*>(a.keys[2].data) : nullptr;
  const uint32_t* pred = reinterpret_cast<const uint32_t*>(f.pred_data);
  uint32_t tk[kTinyGroups];
#pragma unroll
  for (int q = 0; q < kTinyGroups; ++q) tk[q] = 0xFFFFFFFFu;
  uint32_t seen = 0;
  bool lost = false;
  constexpr int kU = 4;   // rows in flight per thread (each brings up to 7 loads)
#pragma unroll 1
  for (int64_t row0 = begin + t; row0 < end; row0 += kU * kTinyThreads) {
    uint32_t key[kU], pv[kU];
    double c[kU][kFusedCols];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t row = min(row0 + u * kTinyThreads, end - 1);   // clamped: the loads need no branch
      pv[u] = pred ? pred[row] : 0u;
      uint32_t k = k0[row];
      if (k1) k |= uint32_t(k1[row]) << 8;
      if (k2) k |= uint32_t(k2[row]) << 16;
      key[u] = k;
#pragma unroll
      for (int j = 0; j < kFusedCols; ++j) c[u][j] = j < f.n_cols ? f.cols[j][row] : 0.0;
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const int64_t row = row0 + u * kTinyThreads;
      if (row < end) {
        bool on = true;
        if (pred) on = f.pred_type == B200_INT32 ? fused_cmp<int32_t>(f.pred_op, int32_t(pv[u]), int32_t(f.pred_scalar))
                                                 : fused_cmp<uint32_t>(f.pred_op, pv[u], uint32_t(f.pred_scalar));
        int g = -1;
#pragma unroll
        for (int q = 0; q < kTinyGroups; ++q)
          if (tk[q] == key[u]) g = q;
        if (g < 0 && on) {   // not in the thread's snapshot: claim the first free entry, or find the key claimed meanwhile
#pragma unroll 1
          for (int q = 0; q < kTinyGroups && g < 0; ++q) {
            const unsigned long long old = atomicCAS(reinterpret_cast<unsigned long long*>(&S.keys[q]), (unsigned long long)kEmpty,
                                                     (unsigned long long)key[u]);
            if (old == kEmpty || old == key[u]) g = q;
          }
#pragma unroll
          for (int q = 0; q < kTinyGroups; ++q) tk[q] = uint32_t(*reinterpret_cast<const volatile uint64_t*>(&S.keys[q]));
          if (g < 0) lost = true;   // a 9th key in this CTA
        }
        if (on && g >= 0 && !((seen >> g) & 1u)) {
          seen |= 1u << g;
          atomicMin(&S.first[g], uint32_t(row - begin));   // rows of a thread come in increasing order
        }
        // No branch around the accumulation: a row the predicate leaves out (or whose key overflowed the table, in
        // which case the call is redone anyway) adds -0.0, the identity of IEEE addition for every accumulator value
        // including -0.0 and +0.0, to an arbitrary group and 0 to its count.  (Lanes that skip the row body make the
        // warp several times slower at a 96 % selection, see gb_tiny8_kernel.)
        const bool add = on && g >= 0;
        const int ga = g >= 0 ? g : 0;
#pragma unroll
        for (int s = 0; s < NS; ++s) {
          double v = c[u][s];   // plain stream s reads column s
          if (s >= f.n_plain) {
            const FusedStream& st = f.streams[s];
            v = fused_arith(st.op1, fused_pick(st.ia, c[u]), fused_arith(st.op2, st.scalar, fused_pick(st.ib, c[u])));
          }
          S.acc[ga][s][t] = __dadd_rn(S.acc[ga][s][t], add ? v : -0.0);
        }
        S.cnt[ga][t] += add ? 1u : 0u;
      }
    }
  }
  if (lost) atomicExch(f.overflow, 1u);
  __syncthreads();
  // ---- totals: warp w reduces the (group, stream) pairs w, w + 8, ... in a fixed order
  for (int pair = warp; pair < kTinyGroups * NS; pair += kTinyThreads / 32) {
    const int g = pair / NS, s = pair % NS;
    double sv = 0.0;
#pragma unroll
    for (int j = 0; j < kTinyThreads / 32; ++j) sv += S.acc[g][s][j * 32 + lane];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) sv += __shfl_down_sync(0xffffffffu, sv, d);
    if (lane == 0) S.totals[g][s] = sv;
  }
  {
    const int g = warp;   // kTinyThreads / 32 == kTinyGroups
    unsigned long long cv = 0;
#pragma unroll
    for (int j = 0; j < kTinyThreads / 32; ++j) cv += S.cnt[g][j * 32 + lane];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) cv += __shfl_down_sync(0xffffffffu, cv, d);
    if (lane == 0) S.counts[g] = cv;
  }
  __syncthreads();
  // ---- the CTA's groups as slot images
  if (t < kTinyGroups) {
    uint64_t* img = reinterpret_cast<uint64_t*>(f.images + (size_t(blockIdx.x) * kTinyGroups + t) * a.s.stride);
    const int words = a.s.stride / 8;
    const uint64_t key = S.keys[t];
    const bool used = key != kEmpty && S.first[t] != 0xFFFFFFFFu;
    img[0] = used ? key : kEmpty;
    img[1] = used ? a.row_base + uint64_t(begin) + S.first[t] : ~uint64_t(0);
    for (int w = 2; w < words; ++w) img[w] = 0;
    if (used) {
      for (int j = 0; j < a.s.n_aggs; ++j) {
        const GbAgg& g = a.s.aggs[j];
        uint64_t* st = img + g.off / 8;
        if (g.kind == B200_HASH_COUNT) {
          st[0] = g.option ? 0 : S.counts[t];   // the argument columns hold no nulls
        } else {
          st[0] = __double_as_longlong(S.totals[t][g.col]);
          st[1] = S.counts[t];
        }
      }
    }
  }
}
