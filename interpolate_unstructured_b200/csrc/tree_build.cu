// tree_build.cu -- the nearest-centre kd-tree built ON THE DEVICE (build_kdtree, m_interp_unstructured.f90:251-262, replaces
// kdtree2_create).  Same tree as the host builder of host_mesh.cpp (kept as IU_B200_HOST_TREE=1 for A/B): complete,
// left-balanced, heap-indexed binary tree over leaf blocks of LEAF centres; an internal node splits along the widest
// extent of the exact bounding box of its centres at the coordinate of its middle element, so that every centre of the
// left child is <= split <= every centre of the right child (what the search's plane bound relies on).
//
// Level-synchronous: the element ranges of all nodes follow from (n_cells, LEAF) alone, so they are tabulated once; per
// level one pass computes the boxes (ordered-integer atomics), one stable radix sort orders all elements by
// (node, coordinate along the node's split dimension) -- a 64-bit sort by coordinate followed by a sort by node id --
// and the middle elements give the splits.  ~10 ms per level for 10M centres, no host work, no copy of the centres.
#include <cub/device/device_radix_sort.cuh>

#include <algorithm>
#include <vector>

#include "device.h"

namespace iu {

namespace {

#define TB_CU(expr)                                  \
  do {                                               \
    cudaError_t e__ = (expr);                        \
    if (e__ != cudaSuccess) return -(int)e__ - 1000; \
  } while (0)

// leaves in the left subtree of a complete (left-balanced) binary tree with l >= 2 leaves (as host_mesh.cpp)
inline int leaves_left(int l) {
  if (l == 2) return 1;
  int d = 0;
  while ((1 << d) < l) ++d;
  return std::min(1 << (d - 1), l - (1 << (d - 2)));
}

// monotone map double -> uint64 (a < b  <=>  enc(a) < enc(b) for non-NaN values; -0.0 sorts just below +0.0)
__device__ __forceinline__ unsigned long long enc(double x) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(x);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ULL);
}
__device__ __forceinline__ double dec(unsigned long long e) {
  const unsigned long long b = (e >> 63) ? (e & 0x7fffffffffffffffULL) : ~e;
  return __longlong_as_double((long long)b);
}

__global__ void k_tb_init(int n, int32_t* perm, int32_t* node_of) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p < n) { perm[p] = p; node_of[p] = 0; }
}

__global__ void k_tb_reset_boxes(int first, int cnt, unsigned long long* mm) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= cnt) return;
  unsigned long long* m = mm + 6 * (size_t)(first + j);
  m[0] = m[1] = m[2] = ~0ULL;   // minima
  m[3] = m[4] = m[5] = 0ULL;    // maxima
}

// exact bounding box of the centres of every internal node of this level (heap indices first .. first+cnt-1)
__global__ void k_tb_boxes(int n, int first, int cnt, const double* __restrict__ centers, const int32_t* __restrict__ perm,
                           const int32_t* __restrict__ node_of, unsigned long long* mm) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const bool in = p < n;
  const int node = in ? node_of[p] : -1;
  const bool act = in && node >= first && node < first + cnt;
  unsigned long long e[3] = {0, 0, 0};
  if (act) {
    const double* c = centers + 3 * (size_t)perm[p];
    e[0] = enc(c[0]); e[1] = enc(c[1]); e[2] = enc(c[2]);
  }
  // positions are grouped by node: most warps lie inside one node and reduce before touching memory
  const unsigned full = 0xffffffffu;
  const int node0 = __shfl_sync(full, node, 0);
  const bool uniform = __all_sync(full, act && node == node0);
  if (uniform) {
    unsigned long long lo[3] = {e[0], e[1], e[2]}, hi[3] = {e[0], e[1], e[2]};
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const unsigned long long a = __shfl_xor_sync(full, lo[k], o), b = __shfl_xor_sync(full, hi[k], o);
        lo[k] = a < lo[k] ? a : lo[k];
        hi[k] = b > hi[k] ? b : hi[k];
      }
    if ((threadIdx.x & 31) == 0) {
      unsigned long long* m = mm + 6 * (size_t)node;
#pragma unroll
      for (int k = 0; k < 3; ++k) { atomicMin(m + k, lo[k]); atomicMax(m + 3 + k, hi[k]); }
    }
  } else if (act) {
    unsigned long long* m = mm + 6 * (size_t)node;
#pragma unroll
    for (int k = 0; k < 3; ++k) { atomicMin(m + k, e[k]); atomicMax(m + 3 + k, e[k]); }
  }
}

// split dimension = widest extent, first maximum (host_mesh.cpp: `if (mx[k] - mn[k] > mx[d] - mn[d]) d = k`)
__global__ void k_tb_dims(int first, int cnt, const unsigned long long* __restrict__ mm, int32_t* __restrict__ sdim) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= cnt) return;
  const unsigned long long* m = mm + 6 * (size_t)(first + j);
  int d = 0;
  double best = dec(m[3]) - dec(m[0]);
  for (int k = 1; k < 3; ++k) {
    const double ext = dec(m[3 + k]) - dec(m[k]);
    if (ext > best) { best = ext; d = k; }
  }
  sdim[first + j] = d;
}

// sort key of every element: its coordinate along the split dimension of the node it is in (elements of nodes that are
// already leaves keep dimension 0: they only have to stay inside their node); value = (node, element)
__global__ void k_tb_keys(int n, int first, int cnt, const double* __restrict__ centers, const int32_t* __restrict__ perm,
                          const int32_t* __restrict__ node_of, const int32_t* __restrict__ sdim,
                          unsigned long long* __restrict__ key, unsigned long long* __restrict__ val) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const int node = node_of[p], el = perm[p];
  const int d = node >= first && node < first + cnt ? sdim[node] : 0;
  key[p] = enc(centers[3 * (size_t)el + d]);
  val[p] = ((unsigned long long)(unsigned)node << 32) | (unsigned)el;
}

__global__ void k_tb_unpack(int n, const unsigned long long* __restrict__ val, uint32_t* __restrict__ node_key, uint32_t* __restrict__ el) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const unsigned long long v = val[p];
  node_key[p] = (uint32_t)(v >> 32);
  el[p] = (uint32_t)v;
}

// split value = the coordinate of the node's middle element: everything left of it is <= it, everything right >= it
__global__ void k_tb_splits(int first, int cnt, const double* __restrict__ centers, const int32_t* __restrict__ perm,
                            const int32_t* __restrict__ lo, const int32_t* __restrict__ sdim, double* __restrict__ split) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= cnt) return;
  const int node = first + j;
  const int mid = lo[2 * node + 2];
  split[node] = centers[3 * (size_t)perm[mid] + sdim[node]];
}

__global__ void k_tb_descend(int n, int first, int cnt, const int32_t* __restrict__ lo, const uint32_t* __restrict__ node_sorted,
                             int32_t* __restrict__ node_of) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const int node = (int)node_sorted[p];
  node_of[p] = node >= first && node < first + cnt ? (p < lo[2 * node + 2] ? 2 * node + 1 : 2 * node + 2) : node;
}

__global__ void k_tb_slots(int n, int n_internal, const int32_t* __restrict__ lo, const int32_t* __restrict__ perm,
                           const int32_t* __restrict__ node_of, int32_t* __restrict__ orig_of_dev) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= n) return;
  const int node = node_of[p];
  orig_of_dev[(size_t)(node - n_internal) * LEAF + (p - lo[node])] = perm[p] + 1;
}

// tight box of a leaf's centres, floats rounded outward (what box_lower_bound assumes)
__global__ void k_tb_leaf_boxes(int n_leaves, int n_internal, const double* __restrict__ centers, const int32_t* __restrict__ perm,
                                const int32_t* __restrict__ lo, const int32_t* __restrict__ hi, float* __restrict__ box) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= n_leaves) return;
  const int node = n_internal + b;
  float* o = box + 8 * (size_t)node;
  double mn[3], mx[3];
  const int a = lo[node], z = hi[node];
  for (int k = 0; k < 3; ++k) mn[k] = mx[k] = centers[3 * (size_t)perm[a] + k];
  for (int p = a + 1; p < z; ++p)
    for (int k = 0; k < 3; ++k) {
      const double v = centers[3 * (size_t)perm[p] + k];
      mn[k] = v < mn[k] ? v : mn[k];
      mx[k] = v > mx[k] ? v : mx[k];
    }
  for (int k = 0; k < 3; ++k) { o[k] = __double2float_rd(mn[k]); o[3 + k] = __double2float_ru(mx[k]); }
  o[6] = 0.0f; o[7] = 0.0f;
}

__global__ void k_tb_union_boxes(int first, int cnt, float* __restrict__ box) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= cnt) return;
  const int i = first + j;
  const float* a = box + 8 * (size_t)(2 * i + 1);
  const float* c = box + 8 * (size_t)(2 * i + 2);
  float* o = box + 8 * (size_t)i;
  for (int k = 0; k < 3; ++k) { o[k] = fminf(a[k], c[k]); o[3 + k] = fmaxf(a[3 + k], c[3 + k]); }
  o[6] = 0.0f; o[7] = 0.0f;
}

// fat nodes {split, left child's split, right child's split, bits(3 x 2-bit split dimension)} (device.h: DeviceGrid::fat)
__global__ void k_tb_fat(int n_internal, const double* __restrict__ split, const int32_t* __restrict__ sdim, double* __restrict__ fat) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_internal) return;
  unsigned long long dims = (unsigned long long)sdim[i];
  double s[2] = {0.0, 0.0};
  for (int c = 0; c < 2; ++c) {
    const long long ch = 2LL * i + 1 + c;
    if (ch < n_internal) { s[c] = split[ch]; dims |= (unsigned long long)sdim[ch] << (2 + 2 * c); }
  }
  fat[4 * (size_t)i] = split[i];
  fat[4 * (size_t)i + 1] = s[0];
  fat[4 * (size_t)i + 2] = s[1];
  fat[4 * (size_t)i + 3] = __longlong_as_double((long long)dims);
}

template <class T>
struct Scratch {
  T* p = nullptr;
  ~Scratch() { if (p) cudaFree(p); }
  cudaError_t alloc(size_t count) { return cudaMalloc((void**)&p, std::max<size_t>(count, 1) * sizeof(T)); }
};

inline int nblk(int64_t n, int t = 256) { return (int)((n + t - 1) / t); }

}  // namespace

// centers: device (3, n_cells).  Outputs, device, allocated by the caller: fat [4 (L-1)] doubles, box [8 (2L-1)] floats,
// orig_of_dev [L LEAF] int32 (1-based original cell of every device slot, 0 = padding), L = ceil(n_cells / LEAF).
// Returns the number of kernels launched (CUB passes counted conservatively), or a negative value on a CUDA error.
int build_kdtree_device(const double* centers, int n_cells, double* fat, float* box, int32_t* orig_of_dev, cudaStream_t s) {
  const int n = n_cells;
  const int L = (n + LEAF - 1) / LEAF;
  if (n <= 0) return 0;
  const int n_internal = L - 1, n_nodes = 2 * L - 1;
  // element range of every node: follows from (n, LEAF) alone
  std::vector<int32_t> h_lo(n_nodes), h_hi(n_nodes), h_l(n_nodes);
  h_lo[0] = 0; h_hi[0] = n; h_l[0] = L;
  for (int i = 0; i < n_internal; ++i) {
    const int ll = leaves_left(h_l[i]);
    const int mid = h_lo[i] + ll * LEAF;
    h_lo[2 * i + 1] = h_lo[i]; h_hi[2 * i + 1] = mid; h_l[2 * i + 1] = ll;
    h_lo[2 * i + 2] = mid; h_hi[2 * i + 2] = h_hi[i]; h_l[2 * i + 2] = h_l[i] - ll;
  }
  Scratch<int32_t> lo, hi, perm, node_of, sdim;
  Scratch<uint32_t> nk_a, nk_b, el_a;
  Scratch<unsigned long long> key_a, key_b, val_a, val_b, mm;
  Scratch<double> split;
  Scratch<char> temp;
  TB_CU(lo.alloc(n_nodes)); TB_CU(hi.alloc(n_nodes)); TB_CU(perm.alloc(n)); TB_CU(node_of.alloc(n)); TB_CU(sdim.alloc(n_internal));
  TB_CU(nk_a.alloc(n)); TB_CU(nk_b.alloc(n)); TB_CU(el_a.alloc(n));
  TB_CU(key_a.alloc(n)); TB_CU(key_b.alloc(n)); TB_CU(val_a.alloc(n)); TB_CU(val_b.alloc(n)); TB_CU(mm.alloc(6 * (size_t)std::max(n_internal, 1)));
  TB_CU(split.alloc(n_internal));
  size_t t1 = 0, t2 = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, t1, key_a.p, key_b.p, val_a.p, val_b.p, n, 0, 64);
  cub::DeviceRadixSort::SortPairs(nullptr, t2, nk_a.p, nk_b.p, el_a.p, (uint32_t*)perm.p, n, 0, 32);
  const size_t temp_bytes = std::max(t1, t2);
  TB_CU(temp.alloc(temp_bytes));
  TB_CU(cudaMemcpyAsync(lo.p, h_lo.data(), (size_t)n_nodes * 4, cudaMemcpyHostToDevice, s));
  TB_CU(cudaMemcpyAsync(hi.p, h_hi.data(), (size_t)n_nodes * 4, cudaMemcpyHostToDevice, s));
  int launches = 0;
  k_tb_init<<<nblk(n), 256, 0, s>>>(n, perm.p, node_of.p); ++launches;
  int depth = 0;
  for (int l = 1; l < L; l <<= 1) ++depth;                      // levels of internal nodes
  for (int lev = 0; lev < depth; ++lev) {
    const int first = (1 << lev) - 1;
    const int cnt = std::min(1 << lev, n_internal - first);
    if (cnt <= 0) break;
    k_tb_reset_boxes<<<nblk(cnt), 256, 0, s>>>(first, cnt, mm.p);
    k_tb_boxes<<<nblk(n), 256, 0, s>>>(n, first, cnt, centers, perm.p, node_of.p, mm.p);
    k_tb_dims<<<nblk(cnt), 256, 0, s>>>(first, cnt, mm.p, sdim.p);
    k_tb_keys<<<nblk(n), 256, 0, s>>>(n, first, cnt, centers, perm.p, node_of.p, sdim.p, key_a.p, val_a.p);
    size_t tb = temp_bytes;
    TB_CU(cub::DeviceRadixSort::SortPairs(temp.p, tb, key_a.p, key_b.p, val_a.p, val_b.p, n, 0, 64, s));
    k_tb_unpack<<<nblk(n), 256, 0, s>>>(n, val_b.p, nk_a.p, el_a.p);
    tb = temp_bytes;
    // stable sort by node id: nodes of this level and the leaves that already ended one level up all have ids < 2^(lev+1)
    TB_CU(cub::DeviceRadixSort::SortPairs(temp.p, tb, nk_a.p, nk_b.p, el_a.p, (uint32_t*)perm.p, n, 0, lev + 2, s));
    k_tb_splits<<<nblk(cnt), 256, 0, s>>>(first, cnt, centers, perm.p, lo.p, sdim.p, split.p);
    k_tb_descend<<<nblk(n), 256, 0, s>>>(n, first, cnt, lo.p, nk_b.p, node_of.p);
    launches += 7 + 2 * 10;
  }
  TB_CU(cudaMemsetAsync(orig_of_dev, 0, (size_t)L * LEAF * 4, s));
  k_tb_slots<<<nblk(n), 256, 0, s>>>(n, n_internal, lo.p, perm.p, node_of.p, orig_of_dev);
  k_tb_leaf_boxes<<<nblk(L), 256, 0, s>>>(L, n_internal, centers, perm.p, lo.p, hi.p, box);
  launches += 2;
  for (int lev = depth - 1; lev >= 0; --lev) {
    const int first = (1 << lev) - 1;
    const int cnt = std::min(1 << lev, n_internal - first);
    if (cnt <= 0) continue;
    k_tb_union_boxes<<<nblk(cnt), 256, 0, s>>>(first, cnt, box); ++launches;
  }
  if (n_internal > 0) { k_tb_fat<<<nblk(n_internal), 256, 0, s>>>(n_internal, split.p, sdim.p, fat); ++launches; }
  TB_CU(cudaGetLastError());
  TB_CU(cudaStreamSynchronize(s));   // the scratch buffers are freed on return
  return launches;
}

}  // namespace iu
