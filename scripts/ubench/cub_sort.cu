// External yardstick, bench-only (never linked into the product library): cub::DeviceRadixSort::SortPairs of the same
// (uint64 key, uint32 row) pairs the library's SortIndices sorts, timed with CUDA events on the same box.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o scripts/ubench/cub_sort scripts/ubench/cub_sort.cu
//   ./cub_sort [rows] [reps]          -> one JSON line
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cub/device/device_radix_sort.cuh>

__global__ void fill(uint64_t* k, uint32_t* v, int64_t n) {
  for (int64_t i = blockIdx.x * int64_t(blockDim.x) + threadIdx.x; i < n; i += int64_t(gridDim.x) * blockDim.x) {
    uint64_t x = uint64_t(i) + 5 * 0x9E3779B97F4A7C15ull;   // splitmix64, the generator family of arrow1_b200/synth.py
    uint64_t z = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    k[i] = z ^ (z >> 31);
    v[i] = uint32_t(i);
  }
}

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); return 2; } } while (0)

int main(int argc, char** argv) {
  const int64_t n = argc > 1 ? atoll(argv[1]) : 100000000;
  const int reps = argc > 2 ? atoi(argv[2]) : 10;
  uint64_t *k_in, *k_out;
  uint32_t *v_in, *v_out;
  CK(cudaMalloc(&k_in, n * 8)); CK(cudaMalloc(&k_out, n * 8));
  CK(cudaMalloc(&v_in, n * 4)); CK(cudaMalloc(&v_out, n * 4));
  fill<<<148 * 8, 256>>>(k_in, v_in, n);
  size_t tmp_bytes = 0;
  CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, k_in, k_out, v_in, v_out, n));
  void* tmp;
  CK(cudaMalloc(&tmp, tmp_bytes));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 3; ++i) CK(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, k_in, k_out, v_in, v_out, n));
  CK(cudaDeviceSynchronize());
  cudaEventRecord(e0);
  for (int i = 0; i < reps; ++i) CK(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, k_in, k_out, v_in, v_out, n));
  cudaEventRecord(e1);
  CK(cudaEventSynchronize(e1));
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  // the same with 32-bit keys (what an int32 / float32 column needs)
  uint32_t *k32_in = reinterpret_cast<uint32_t*>(k_in), *k32_out = reinterpret_cast<uint32_t*>(k_out);
  size_t tmp32 = 0;
  CK(cub::DeviceRadixSort::SortPairs(nullptr, tmp32, k32_in, k32_out, v_in, v_out, n));
  float ms32 = -1;
  if (tmp32 <= tmp_bytes) {
    for (int i = 0; i < 3; ++i) CK(cub::DeviceRadixSort::SortPairs(tmp, tmp32, k32_in, k32_out, v_in, v_out, n));
    cudaEventRecord(e0);
    for (int i = 0; i < reps; ++i) CK(cub::DeviceRadixSort::SortPairs(tmp, tmp32, k32_in, k32_out, v_in, v_out, n));
    cudaEventRecord(e1);
    CK(cudaEventSynchronize(e1));
    cudaEventElapsedTime(&ms32, e0, e1);
    ms32 /= reps;
  }
  printf("{\"yardstick\": \"cub::DeviceRadixSort::SortPairs\", \"cub_version\": %d, \"rows\": %lld, \"reps\": %d, "
         "\"u64_key_u32_value_ms\": %.4f, \"u64_rows_per_s\": %.4g, \"u32_key_u32_value_ms\": %.4f, \"temp_bytes\": %zu}\n",
         CUB_VERSION, (long long)n, reps, ms / reps, double(n) / (ms / reps * 1e-3), ms32, tmp_bytes);
  return 0;
}
