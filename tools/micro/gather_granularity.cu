// DRAM bytes fetched per scattered 4-byte / 24-byte read on B200, by load flavour (run under ncu, dram__bytes_read.sum):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_granularity gather_granularity.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t mix(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL; z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL; return z ^ (z >> 31);
}
template <int FLAVOUR>
__device__ __forceinline__ int ld4(const int* p) {
  int v;
  if (FLAVOUR == 0) asm volatile("ld.global.b32 %0, [%1];" : "=r"(v) : "l"(p));
  else if (FLAVOUR == 1) asm volatile("ld.global.nc.b32 %0, [%1];" : "=r"(v) : "l"(p));
  else if (FLAVOUR == 2) asm volatile("ld.global.cs.b32 %0, [%1];" : "=r"(v) : "l"(p));
  else if (FLAVOUR == 3) asm volatile("ld.global.cg.b32 %0, [%1];" : "=r"(v) : "l"(p));
  else if (FLAVOUR == 4) asm volatile("ld.global.L1::no_allocate.b32 %0, [%1];" : "=r"(v) : "l"(p));
  else if (FLAVOUR == 5) asm volatile("ld.global.L1::no_allocate.L2::64B.b32 %0, [%1];" : "=r"(v) : "l"(p));
  else if (FLAVOUR == 6) asm volatile("ld.global.lu.b32 %0, [%1];" : "=r"(v) : "l"(p));
  else asm volatile("ld.volatile.global.b32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
template <int FLAVOUR>
__global__ void k_gather4(const int* __restrict__ a, size_t n, size_t m, int* out) {
  int acc = 0;
  for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < m; t += (size_t)gridDim.x * blockDim.x)
    acc += ld4<FLAVOUR>(a + mix(t * 0x9E3779B97F4A7C15ULL) % n);
  if (acc == 0x7fffffff) *out = acc;
}
// one random 4-byte read per 32-byte sector-sized stride / per 128-byte line (each line touched exactly once)
__global__ void k_stride(const int* __restrict__ a, size_t n_lines, int stride_ints, int* out) {
  int acc = 0;
  for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < n_lines; t += (size_t)gridDim.x * blockDim.x) {
    const size_t line = mix(t) % n_lines;
    acc += ld4<2>(a + line * stride_ints);
  }
  if (acc == 0x7fffffff) *out = acc;
}
int main() {
  const size_t n = (size_t)1 << 30;   // 4 GB of ints
  const size_t m = (size_t)1 << 26;   // 67M reads
  int *a, *out;
  cudaMalloc(&a, n * 4); cudaMalloc(&out, 4);
  cudaMemset(a, 0, n * 4);
  k_gather4<0><<<148 * 8, 256>>>(a, n, m, out);
  k_gather4<1><<<148 * 8, 256>>>(a, n, m, out);
  k_gather4<2><<<148 * 8, 256>>>(a, n, m, out);
  k_gather4<3><<<148 * 8, 256>>>(a, n, m, out);
  k_gather4<4><<<148 * 8, 256>>>(a, n, m, out);
  k_gather4<5><<<148 * 8, 256>>>(a, n, m, out);
  k_gather4<6><<<148 * 8, 256>>>(a, n, m, out);
  k_gather4<7><<<148 * 8, 256>>>(a, n, m, out);
  k_stride<<<148 * 8, 256>>>(a, m / 2, 32, out);   // every touched 128-byte line is distinct
  k_stride<<<148 * 8, 256>>>(a, m / 2, 16, out);   // 64-byte stride
  k_stride<<<148 * 8, 256>>>(a, m / 2, 8, out);    // 32-byte stride
  cudaDeviceSynchronize();
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
