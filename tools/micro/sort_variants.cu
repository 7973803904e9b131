// Radix-sort variants for the coherence sort of 2^26 (key, index) pairs with 24-bit keys on B200:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o sort_variants sort_variants.cu
#include <cstdio>
#include <cstdint>
#include <cub/device/device_radix_sort.cuh>
#include <cuda_runtime.h>

__global__ void k_fill(uint32_t* k, uint32_t* v, uint64_t* kv, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    uint64_t z = (i + 12345) * 0x9E3779B97F4A7C15ULL; z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL; z ^= z >> 31;
    const uint32_t key = (uint32_t)(z & 0xffffffu);
    k[i] = key; v[i] = (uint32_t)i; kv[i] = ((uint64_t)key << 32) | (uint32_t)i;
  }
}
template <class F> float timed(F f, int reps = 5) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(a); for (int i = 0; i < reps; ++i) f(); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b); return ms / reps;
}
int main() {
  const size_t n = (size_t)1 << 26;
  uint32_t *k, *k2, *v, *v2; uint64_t *kv, *kv2; void* tmp; size_t tb = 0, tb2 = 0;
  cudaMalloc(&k, n * 4); cudaMalloc(&k2, n * 4); cudaMalloc(&v, n * 4); cudaMalloc(&v2, n * 4); cudaMalloc(&kv, n * 8); cudaMalloc(&kv2, n * 8);
  k_fill<<<148 * 8, 256>>>(k, v, kv, n);
  cub::DeviceRadixSort::SortPairs(nullptr, tb, k, k2, v, v2, n, 0, 24);
  cub::DeviceRadixSort::SortKeys(nullptr, tb2, kv, kv2, n, 32, 56);
  if (tb2 > tb) tb = tb2;
  cudaMalloc(&tmp, tb);
  printf("pairs u32/u32, bits 0..24      : %.3f ms\n", timed([&] { cub::DeviceRadixSort::SortPairs(tmp, tb, k, k2, v, v2, n, 0, 24); }));
  printf("pairs u32/u32, bits 0..32      : %.3f ms\n", timed([&] { cub::DeviceRadixSort::SortPairs(tmp, tb, k, k2, v, v2, n, 0, 32); }));
  printf("pairs u32/u32, bits 0..16      : %.3f ms\n", timed([&] { cub::DeviceRadixSort::SortPairs(tmp, tb, k, k2, v, v2, n, 0, 16); }));
  printf("keys u64 (key<<32|idx), 32..56 : %.3f ms\n", timed([&] { cub::DeviceRadixSort::SortKeys(tmp, tb, kv, kv2, n, 32, 56); }));
  cub::DoubleBuffer<uint32_t> dk(k, k2), dv(v, v2);
  printf("pairs DoubleBuffer, bits 0..24 : %.3f ms\n", timed([&] { cub::DeviceRadixSort::SortPairs(tmp, tb, dk, dv, n, 0, 24); }));
  printf("keys only u32, bits 0..24      : %.3f ms\n", timed([&] { cub::DeviceRadixSort::SortKeys(tmp, tb, k, k2, n, 0, 24); }));
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
