// Microbenchmark (not product code): shared-memory accumulation throughput on sm_100a for the
// group-by kernel design.  Each thread performs ITERS updates at pseudo-random slots of a G-slot table.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t lcg(uint32_t& s) { s = s * 1664525u + 1013904223u; return s >> 8; }

template <int MODE>
__global__ void __launch_bounds__(512) bench(int G, int iters, double* sink, int warps_private) {
  extern __shared__ __align__(16) unsigned char smem[];
  double* dsum = reinterpret_cast<double*>(smem);
  unsigned long long* usum = reinterpret_cast<unsigned long long*>(smem);
  uint32_t* cnt = reinterpret_cast<uint32_t*>(smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total_words = (MODE == 3 || MODE == 4) ? G * 2 * warps_private + G : G * 2;
  for (int i = threadIdx.x; i < total_words; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  __syncthreads();
  uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
  double v = 1.0 + threadIdx.x * 1e-3;
  if (MODE == 0) {            // native u32 add, no return
    for (int i = 0; i < iters; ++i) atomicAdd(&cnt[lcg(s) % G], 1u);
  } else if (MODE == 1) {     // f64 atomicAdd (CAS spin)
    for (int i = 0; i < iters; ++i) atomicAdd(&dsum[lcg(s) % G], v);
  } else if (MODE == 2) {     // u64 atomicAdd (CAS spin)
    for (int i = 0; i < iters; ++i) atomicAdd(&usum[lcg(s) % G], (unsigned long long)i);
  } else if (MODE == 3) {     // warp-private table, tag election in smem, plain RMW
    double* mine = dsum + (warp % warps_private) * G;
    uint8_t* tag = reinterpret_cast<uint8_t*>(dsum + warps_private * G) + (warp % warps_private) * G;  // needs G*wp bytes
    for (int i = 0; i < iters; ++i) {
      const uint32_t g = lcg(s) % G;
      bool done = false;
      while (true) {
        if (!done) tag[g] = (uint8_t)lane;
        __syncwarp();
        if (!done && tag[g] == lane) { mine[g] += v; done = true; }
        __syncwarp();
        if (__all_sync(0xffffffffu, done)) break;
      }
    }
  } else if (MODE == 4) {     // warp-private table, match_any election, plain RMW by leaders w/ shuffle-free retry
    double* mine = dsum + (warp % warps_private) * G;
    for (int i = 0; i < iters; ++i) {
      const uint32_t g = lcg(s) % G;
      unsigned m = __match_any_sync(0xffffffffu, g);
      // serialize duplicates: lanes take turns by rank inside their match group
      int rank = __popc(m & ((1u << lane) - 1));
      int maxr = __reduce_max_sync(0xffffffffu, rank);
      for (int r = 0; r <= maxr; ++r) {
        if (rank == r) mine[g] += v;
        __syncwarp();
      }
    }
  } else if (MODE == 5) {     // native u32 add WITH return
    uint32_t acc = 0;
    for (int i = 0; i < iters; ++i) acc += atomicAdd(&cnt[lcg(s) % G], 1u);
    if (acc == 0xdeadbeef) sink[0] = acc;
  } else if (MODE == 6) {     // plain LDS/STS RMW (racy, upper bound)
    for (int i = 0; i < iters; ++i) { const uint32_t g = lcg(s) % G; dsum[g] += v; }
  } else if (MODE == 7) {     // checked min/max: LDS.128 + compare, rare CAS
    unsigned long long* mm = usum;
    for (int i = 0; i < iters; ++i) {
      const uint32_t g = lcg(s) % (G / 2);
      const unsigned long long x = (unsigned long long)lcg(s) << 20;
      ulonglong2 cur = *reinterpret_cast<ulonglong2*>(&mm[2 * g]);
      if (x > cur.y) atomicMax(&mm[2 * g + 1], x);
      if (~x > cur.x) atomicMax(&mm[2 * g], ~x);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) sink[blockIdx.x] = dsum[0];
}

template <int MODE>
void run(const char* name, int G, int threads, int wp) {
  int iters = 4096;
  double* sink; cudaMalloc(&sink, 148 * 8 * sizeof(double));
  size_t smem = (MODE == 3 || MODE == 4) ? size_t(G) * 8 * wp + size_t(G) * wp + 64 : size_t(G) * 8;
  cudaFuncSetAttribute(bench<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  bench<MODE><<<148, threads, smem>>>(G, 64, sink, wp);
  cudaEventRecord(e0);
  bench<MODE><<<148, threads, smem>>>(G, iters, sink, wp);
  cudaEventRecord(e1); cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  cudaError_t err = cudaGetLastError();
  double ops = double(148) * threads * iters;
  printf("%-28s G=%5d threads=%4d: %8.3f ms  %.2f updates/ns  %.2f updates/cycle/SM(@1.9GHz) %s\n", name, G, threads, ms,
         ops / ms / 1e6, ops / ms / 1e6 / 148 / 1.9, err == cudaSuccess ? "" : cudaGetErrorString(err));
  cudaFree(sink);
}

int main() {
  for (int G : {4, 64, 1024}) {
    for (int threads : {256, 512}) {
      run<0>("u32 ATOMS.ADD", G, threads, 1);
      run<5>("u32 ATOMS.ADD +return", G, threads, 1);
      run<1>("f64 atomicAdd (CAS)", G, threads, 1);
      run<2>("u64 atomicAdd (CAS)", G, threads, 1);
      run<3>("warp-private tag election", G, threads, threads / 32);
      run<4>("warp-private match_any", G, threads, threads / 32);
      run<6>("plain RMW (racy bound)", G, threads, 1);
      run<7>("checked min/max", G, threads, 1);
    }
  }
  return 0;
}
