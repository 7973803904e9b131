// kernels.cu -- sm_100a kernels of libiu_b200.so.
//
// Compiled with --fmad=false: every fp64 expression below evaluates in source order with one
// rounding per operation, exactly like the reference's `gfortran -O2` x86-64 (SSE2, no FMA)
// build (Makefile:7-8).  fp64 '/' and sqrt() are IEEE round-to-nearest on the device.  Branch
// decisions (face choice, walk termination, nearest centre) are therefore bit-identical to the
// reference arithmetic restated in oracle/iu_oracle.cpp.
//
// Kernels:
//   k_cell_centers     iu_get_cell_center for every cell (input of the kd-tree build)
//   k_build_records    set_cell_points + set_face_normal_vectors + set_cell_volumes -> 256/192-B
//                      per-cell records in kd-tree leaf order, plus the leaf centre blocks
//   k_morton_keys      sort key of each query for the coherence sort: position of its key cell along a Hilbert curve
//                      (3-D meshes) or the Z curve (planar meshes); cub radix sort follows
//   k_query<...>       THE hot kernel: [kd-tree nearest centre | hint] -> ray walk -> interpolate
//   k_walk<...>        iu_get_cell_through_neighbors with explicit r0 / ic0 / mask outputs
//   k_ray_length_sample  mean ray length of a batch (decides whether k_walk processes it in sorted order)
//   k_trace<...>       iu_integrate_along_field, one field line per thread
//   k_pack_point_data, k_export_geometry, k_generate_queries: plumbing
#include <cub/device/device_radix_sort.cuh>

#include <cstdlib>
#include <string>

#include "device.h"

namespace iu {

namespace {

constexpr double TINY_DISTANCE = 1e-100;   // m_interp_unstructured.f90:20
constexpr double HUGE_DISTANCE = 1e100;    // :738
// -DIU_CHECKED=1 (make LIBDIR=../lib_checked EXTRA=-DIU_CHECKED=1): every index the kernels derive from mesh or query
// data (permutation entries, start / neighbour cells, leaf slots, voxel ids, point ids, original cell ids) is range
// checked before it is used; violations are counted (IU_B200_STAT_BOUNDS_ERRORS) and the access is skipped.  This is the
// repository's own memcheck: compute-sanitizer is closed on the GPU pool this was developed on (tools/sanitize_smoke.py).
#ifndef IU_CHECKED
#define IU_CHECKED 0
#endif
#if IU_CHECKED
__device__ unsigned long long g_bounds_errors = 0;
__device__ long long g_lim[4] = {0, 0, 0, 0};   // device slots, points, voxels, leaves: set before every launch
#define IU_OOB(cond) ((cond) ? (atomicAdd(&g_bounds_errors, 1ULL), true) : false)
#else
#define IU_OOB(cond) false
#endif
#ifndef IU_QUERY_THREADS
#define IU_QUERY_THREADS 256
#endif
constexpr int QUERY_THREADS = IU_QUERY_THREADS;
#ifndef IU_QUERY_MIN_BLOCKS
#define IU_QUERY_MIN_BLOCKS 3   // resident CTAs per SM the hot kernel is compiled for: 3 = 80 registers, no spills.  With the
                                // round-1 tree search 4 CTAs at 64 registers (64 B of spills) were faster; with the voxel-table
                                // search 3 are: C3 +1.9 %, C5 +6 %, C4 +3 %, C2 +2 % (profiles/experiments/r03b_*)
#endif
#ifndef IU_SELP
#define IU_SELP 1          // 1: branch-free selection of the query coordinate of a node's split dimension
#endif

#ifndef IU_WALK_256
#define IU_WALK_256 0   // 1: 32-byte loads in the walk too (6 instead of 12 per cell; spills at 64 and at 80 registers: slower both times, profiles/experiments/r03d_*)
#endif
#ifndef IU_HILBERT
#define IU_HILBERT 1    // 1: queries of 3-D meshes are ordered along a Hilbert curve, 0: Morton order
#endif
#ifndef IU_NODE_BOX
#define IU_NODE_BOX 1   // 1: test the far child's tight centre box before descending into it (nearest_center)
#endif
// Query inputs and results are touched exactly once per launch: mark them evict-first (ld/st .cs)
// so that they do not push the cell records of the current window of the sort order out of L2.
#ifndef IU_STREAM_HINTS
#define IU_STREAM_HINTS 1
#endif
// A load that misses L2 fetches a whole 128-byte line from DRAM by default on B200 (tools/micro/gather_granularity.cu:
// 125 bytes of DRAM reads per scattered 4-byte load); the .L2::64B qualifier halves that (63 bytes).  The scattered
// query IO (24-byte coordinates, 4-byte hints through the sort permutation) asks for 64 bytes; IU_LTC64_MESH does the
// same for the 32 / 64-byte gathers of the search structures and the point data.
#ifndef IU_LTC64
#define IU_LTC64 1
#endif
#ifndef IU_LTC64_MESH
#define IU_LTC64_MESH 0
#endif
#if IU_LTC64
#define IU_LTC_Q ".L2::64B"
#else
#define IU_LTC_Q ""
#endif
#if IU_LTC64_MESH
#define IU_LTC_M ".L2::64B"
#else
#define IU_LTC_M ""
#endif
#if IU_STREAM_HINTS
#define IU_CS ".cs"
#define ST_STREAM(p, v) __stcs(p, v)
#else
#define IU_CS ""
#define ST_STREAM(p, v) (*(p) = (v))
#endif
__device__ __forceinline__ double ld_stream(const double* p) {
  double v; asm volatile("ld.global" IU_CS IU_LTC_Q ".f64 %0, [%1];" : "=d"(v) : "l"(p)); return v;
}
__device__ __forceinline__ double2 ld_stream(const double2* p) {
  double2 v; asm volatile("ld.global" IU_CS IU_LTC_Q ".v2.f64 {%0,%1}, [%2];" : "=d"(v.x), "=d"(v.y) : "l"(p)); return v;
}
__device__ __forceinline__ int ld_stream(const int* p) {
  int v; asm volatile("ld.global" IU_CS IU_LTC_Q ".b32 %0, [%1];" : "=r"(v) : "l"(p)); return v;
}
__device__ __forceinline__ uint32_t ld_stream(const uint32_t* p) {
  uint32_t v; asm volatile("ld.global" IU_CS IU_LTC_Q ".b32 %0, [%1];" : "=r"(v) : "l"(p)); return v;
}
#define LD_STREAM(p) ld_stream(p)

struct V3 { double x, y, z; };
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 neg(V3 a) { return {-a.x, -a.y, -a.z}; }
// Fortran dot_product: ((a1*b1 + a2*b2) + a3*b3)
__device__ __forceinline__ double dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
// cross_product, :644-651
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
__device__ __forceinline__ double cross_z(V3 a, V3 b) { return a.x * b.y - a.y * b.x; }   // :654-657
__device__ __forceinline__ double stp(V3 a, V3 b, V3 c) { return dot3(a, cross(b, c)); }    // :659-662

// gfortran NORM2: running-scale algorithm with scale starting at 1.  While all
// |x_i| <= 1 the scale never changes and val = |x_i| / 1 is exact, so the result is
// sqrt(x1^2 + x2^2 + x3^2) summed in order: that is the fast path.
__device__ __forceinline__ double norm2_3(V3 v) {
  const double ax = fabs(v.x), ay = fabs(v.y), az = fabs(v.z);
  if (ax <= 1.0 && ay <= 1.0 && az <= 1.0) return sqrt((ax * ax + ay * ay) + az * az);
  double scale = 1.0, s = 0.0;
  const double a[3] = {ax, ay, az};
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    if (a[i] != 0.0) {
      if (scale < a[i]) {
        const double val = scale / a[i];
        s = 1.0 + s * val * val;
        scale = a[i];
      } else {
        const double val = a[i] / scale;
        s += val * val;
      }
    }
  }
  return scale * sqrt(s);
}

__device__ __forceinline__ V3 ld3(const double* p) { return {__ldg(p), __ldg(p + 1), __ldg(p + 2)}; }

// 32-byte load (sm_100a: LDG.E.256): one L1 tag lookup per lane for a whole leaf slot / a third of
// a record's normals instead of two.  p must be 32-byte aligned.
struct D4 { double x, y, z, w; };
#ifndef IU_LD256_SPLIT
#define IU_LD256_SPLIT 0   // 1: every 32-byte gather as two 16-byte loads (A/B of the LSU data-pipe cost of LDG.256)
#endif
__device__ __forceinline__ D4 ld256(const double* p) {
  D4 r;
#if IU_LD256_SPLIT
  asm volatile("ld.global.nc" IU_LTC_M ".v2.f64 {%0,%1}, [%2];" : "=d"(r.x), "=d"(r.y) : "l"(p));
  asm volatile("ld.global.nc" IU_LTC_M ".v2.f64 {%0,%1}, [%2];" : "=d"(r.z), "=d"(r.w) : "l"(p + 2));
#else
  asm volatile("ld.global.nc" IU_LTC_M ".v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w) : "l"(p));
#endif
  return r;
}
__device__ __forceinline__ uint4 ld128u(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc" IU_LTC_M ".v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
// 4 consecutive (x, y, z) triples at a 32-byte aligned address as three 32-byte loads
__device__ __forceinline__ void ld3x4(const double* p, V3* out) {
  const D4 a = ld256(p), b = ld256(p + 4), c = ld256(p + 8);
  out[0] = V3{a.x, a.y, a.z}; out[1] = V3{a.w, b.x, b.y}; out[2] = V3{b.z, b.w, c.x}; out[3] = V3{c.y, c.z, c.w};
}

// PPC consecutive (x, y, z) triples starting at the 16-byte aligned address p, fetched as 16-byte
// loads (PPC = 3 reads one trailing double that belongs to the record: area / pad)
template <int PPC>
__device__ __forceinline__ void ld3n(const double* p, V3* out) {
  constexpr int N2 = (3 * PPC + 1) / 2;
  double d[2 * N2];
  const double2* p2 = reinterpret_cast<const double2*>(p);
#pragma unroll
  for (int j = 0; j < N2; ++j) { const double2 v = __ldg(p2 + j); d[2 * j] = v.x; d[2 * j + 1] = v.y; }
#pragma unroll
  for (int k = 0; k < PPC; ++k) out[k] = V3{d[3 * k], d[3 * k + 1], d[3 * k + 2]};
}

// ------------------------------------------------------------------------------------------------
// set-up kernels
// ------------------------------------------------------------------------------------------------
template <int PPC>
__device__ __forceinline__ V3 center_of(const V3* p) {
  V3 s = p[0];
#pragma unroll
  for (int k = 1; k < PPC; ++k) s = s + p[k];
  const double d = (double)PPC;
  return {s.x / d, s.y / d, s.z / d};   // sum(cell_points, dim=2) / n_points_per_cell, :268
}

template <int PPC>
__global__ void k_cell_centers(int n_cells, const double* __restrict__ points, const int32_t* __restrict__ cells,
                               double* __restrict__ centers) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_cells) return;
  V3 p[PPC];
#pragma unroll
  for (int k = 0; k < PPC; ++k) p[k] = ld3(points + 3 * (size_t)(cells[(size_t)c * PPC + k] - 1));
  const V3 ce = center_of<PPC>(p);
  centers[3 * (size_t)c] = ce.x; centers[3 * (size_t)c + 1] = ce.y; centers[3 * (size_t)c + 2] = ce.z;
}

__global__ void k_fill_dev_of_orig(int n_slots, const int32_t* __restrict__ orig_of_dev, int32_t* __restrict__ dev_of_orig) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_slots) return;
  const int o = orig_of_dev[s];
  if (o > 0) dev_of_orig[o] = s + 1;
}

// One thread per device slot: gathers the cell's vertices and derives normals / volume with the
// reference's arithmetic (set_cell_points :291-302, set_face_normal_vectors :306-370,
// set_cell_volumes :372-410).
template <int PPC, int CT>
__global__ void k_build_records(int n_slots, const int32_t* __restrict__ orig_of_dev, const int32_t* __restrict__ dev_of_orig,
                                const double* __restrict__ points, const int32_t* __restrict__ cells,
                                const int32_t* __restrict__ neighbors, const double* __restrict__ centers,
                                double* __restrict__ rec, double* __restrict__ centers4, double* __restrict__ cen32) {
  using L = RecLayout<PPC>;
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_slots) return;
  const int o = orig_of_dev[s];
  double* R = rec + (size_t)s * L::STRIDE;
  double* C3 = centers4 + 24 * (size_t)(s / LEAF) + (s % LEAF);   // leaf block: x[8], y[8], z[8]
  double* C4 = cen32 + 4 * (size_t)s;                             // {x, y, z, original id}
  if (o <= 0) {   // padding slot of the last leaf block: never the nearest, never visited
    const double inf = __longlong_as_double(0x7ff0000000000000LL);
    C3[0] = inf; C3[LEAF] = inf; C3[2 * LEAF] = inf;
    C4[0] = inf; C4[1] = inf; C4[2] = inf; C4[3] = 0.0;
    for (int i = 0; i < L::STRIDE; ++i) R[i] = 0.0;
    return;
  }
  const size_t c = (size_t)(o - 1);
  V3 p[PPC];
  int32_t pid[PPC];
#pragma unroll
  for (int k = 0; k < PPC; ++k) {
    pid[k] = cells[c * PPC + k] - 1;
    p[k] = ld3(points + 3 * (size_t)pid[k]);
  }
  const V3 center = center_of<PPC>(p);
  V3 normal_cell{0, 0, 0};
  if (CT != IU_TETRA_CT) normal_cell = cross(p[1] - p[0], p[2] - p[1]);          // :322-324
#pragma unroll
  for (int k = 0; k < PPC; ++k) {
    const int k1 = (k + 1) % PPC, k2 = (k + 2) % PPC;
    V3 nf;
    if (CT != IU_TETRA_CT) nf = cross(p[k1] - p[k], normal_cell);               // :329-330
    else nf = cross(p[k1] - p[k], p[k2] - p[k1]);                               // :351-353
    if (dot3(p[k] - center, nf) < 0) nf = neg(nf);                              // :333-334, :356-357
    const double len = norm2_3(nf);
    R[L::NRM + 3 * k] = nf.x / len; R[L::NRM + 3 * k + 1] = nf.y / len; R[L::NRM + 3 * k + 2] = nf.z / len;
    R[L::PTS + 3 * k] = p[k].x; R[L::PTS + 3 * k + 1] = p[k].y; R[L::PTS + 3 * k + 2] = p[k].z;
  }
  double vol;
  if (CT == IU_TRIANGLE_CT) vol = 0.5 * norm2_3(cross(p[1] - p[0], p[2] - p[0]));
  else if (CT == IU_QUAD_CT) {
    const double a1 = 0.5 * norm2_3(cross(p[1] - p[0], p[2] - p[0]));
    const double a2 = 0.5 * norm2_3(cross(p[2] - p[0], p[PPC - 1] - p[0]));
    vol = a1 + a2;
  } else {
    const double sixth = 1 / 6.0;
    vol = sixth * stp(p[1] - p[0], p[2] - p[0], p[PPC - 1] - p[0]);
  }
  R[L::VOL] = vol;
  int32_t* RI = reinterpret_cast<int32_t*>(R);
#pragma unroll
  for (int k = 0; k < PPC; ++k) {
    const int nb = neighbors[c * PPC + k];
    RI[L::NBR_I + k] = nb >= 1 ? dev_of_orig[nb] : nb;
    RI[L::CEL_I + k] = pid[k];
  }
  RI[L::ORG_I] = o;
  if (PPC == 4) { RI[L::ORG_I + 1] = 0; R[30] = 0.0; R[31] = 0.0; }   // [27] hi word, [30], [31]: padding
  else { R[19] = 0.0; RI[L::CEL_I + 3] = 0; }
  // leaf centre block entry; the centre is the array the tree was built from (bit-identical to
  // sum(cell_points)/ppc at :429)
  C3[0] = centers[3 * c]; C3[LEAF] = centers[3 * c + 1]; C3[2 * LEAF] = centers[3 * c + 2];
  C4[0] = centers[3 * c]; C4[1] = centers[3 * c + 1]; C4[2] = centers[3 * c + 2];
  C4[3] = __hiloint2double(0, o);
}

template <int PPC>
__global__ void k_export_geometry(int n_slots, const int32_t* __restrict__ orig_of_dev, const double* __restrict__ rec,
                                  const double* __restrict__ centers4, double* cell_points, double* normals,
                                  double* volume, double* centers) {
  using L = RecLayout<PPC>;
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_slots) return;
  const int o = orig_of_dev[s];
  if (o <= 0) return;
  const double* R = rec + (size_t)s * L::STRIDE;
  const size_t c = (size_t)(o - 1);
  for (int i = 0; i < 3 * PPC; ++i) {
    if (cell_points) cell_points[c * 3 * PPC + i] = R[L::PTS + i];
    if (normals) normals[c * 3 * PPC + i] = R[L::NRM + i];
  }
  if (volume) volume[c] = R[L::VOL];
  if (centers) for (int i = 0; i < 3; ++i) centers[3 * c + i] = centers4[24 * (size_t)(s / LEAF) + i * LEAF + (s % LEAF)];
}

// the raw mesh back from the records: cells(ppc, Nc) 1-based in original order, points(3, Np)
// (every vertex of a cell carries its coordinates; vertices shared by several cells are written
// several times with the same bits)
template <int PPC>
__global__ void k_export_mesh(int n_slots, const int32_t* __restrict__ orig_of_dev, const double* __restrict__ rec,
                              double* points, int32_t* cells) {
  using L = RecLayout<PPC>;
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= n_slots) return;
  const int o = orig_of_dev[s];
  if (o <= 0) return;
  const double* R = rec + (size_t)s * L::STRIDE;
  const int32_t* RI = reinterpret_cast<const int32_t*>(R);
  for (int k = 0; k < PPC; ++k) {
    const int pid = RI[L::CEL_I + k];
    if (cells) cells[(size_t)(o - 1) * PPC + k] = pid + 1;
    if (points) for (int d = 0; d < 3; ++d) points[3 * (size_t)pid + d] = R[L::PTS + 3 * k + d];
  }
}

// column-major host layout (n_points, n_cols) -> point-major device table
__global__ void k_pack_point_data(int n_points, int n_cols, int col0, int stride, const double* __restrict__ cols,
                                  double* __restrict__ pd) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)n_points * n_cols) return;
  const int v = (int)(i / n_points);
  const size_t p = i % n_points;
  pd[p * stride + col0 + v] = cols[i];
}

// the reverse of k_pack_point_data
__global__ void k_unpack_point_data(int n_points, int n_cols, int col0, int stride, const double* __restrict__ pd,
                                    double* __restrict__ cols) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)n_points * n_cols) return;
  const int v = (int)(i / n_points);
  const size_t p = i % n_points;
  cols[i] = pd[p * stride + col0 + v];
}

__global__ void k_repack_point_data(int n_points, int n_cols, int old_stride, int new_stride, const double* __restrict__ a,
                                    double* __restrict__ b) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)n_points * new_stride) return;
  const size_t p = i / new_stride;
  const int v = (int)(i % new_stride);
  b[i] = v < n_cols ? a[p * old_stride + v] : 0.0;
}

__device__ __forceinline__ uint64_t splitmix64(uint64_t idx, uint64_t seed) {
  uint64_t z = (idx + seed) * 0x9E3779B97F4A7C15ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}

__global__ void k_generate_queries(int64_t n, int64_t first, uint64_t seed, double lox, double loy, double loz, double hix,
                                   double hiy, double hiz, double* __restrict__ r) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 3 * n) return;
  const int k = (int)(i % 3);
  const double u = (double)(splitmix64((uint64_t)(i + 3 * first), seed) >> 11) * (1.0 / 9007199254740992.0);
  const double lo = k == 0 ? lox : k == 1 ? loy : loz;
  const double hi = k == 0 ? hix : k == 1 ? hiy : hiz;
  r[i] = lo + u * (hi - lo);
}

// dst row i <- src row i where i_cell[i] >= 1 (rows of not-found queries stay untouched)
__global__ void k_merge_rows(int64_t n, int row_words, const int32_t* __restrict__ i_cell, const uint32_t* __restrict__ src,
                             uint32_t* __restrict__ dst) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n * row_words) return;
  if (i_cell[i / row_words] >= 1) dst[i] = src[i];
}

// ------------------------------------------------------------------------------------------------
// Morton keys
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t spread3(uint32_t v) {   // 10 bits -> every third bit
  v &= 0x3ffu;
  v = (v | (v << 16)) & 0x030000ffu;
  v = (v | (v << 8)) & 0x0300f00fu;
  v = (v | (v << 4)) & 0x030c30c3u;
  v = (v | (v << 2)) & 0x09249249u;
  return v;
}
__device__ __forceinline__ uint32_t spread2(uint32_t v) {   // 16 bits -> every second bit
  v &= 0xffffu;
  v = (v | (v << 8)) & 0x00ff00ffu;
  v = (v | (v << 4)) & 0x0f0f0f0fu;
  v = (v | (v << 2)) & 0x33333333u;
  v = (v | (v << 1)) & 0x55555555u;
  return v;
}
__device__ __forceinline__ uint32_t quantise(double x, double lo, double scale, uint32_t maxv) {
  const double t = (x - lo) * scale;
  if (!(t > 0.0)) return 0u;   // also NaN
  return t >= (double)maxv ? maxv : (uint32_t)t;
}

// One level of the 3-D Hilbert curve as a table: entry [state * 8 + octant] = digit | next state << 3, where
// octant = x | y << 1 | z << 2 bits of that level and state = 3 * e + d of the compact-Hilbert-index recurrence
// (entry point e in 0..7, direction d in 0..2): 24 states x 8 octants.  Any continuous space-filling order would do -
// the order only decides which queries share a warp and an L2 window, never a result.
__device__ __forceinline__ uint32_t hilbert_table_entry(int state, int octant) {
  auto rotr = [](int x, int i) { i %= 3; return ((x >> i) | (x << (3 - i))) & 7; };
  auto rotl = [](int x, int i) { i %= 3; return ((x << i) | (x >> (3 - i))) & 7; };
  auto trailing_ones = [](int x) { return __ffs(~x) - 1; };
  const int e = state / 3, d = state % 3;
  const int l = rotr(octant ^ e, d + 1);
  const int w = l ^ (l >> 1) ^ (l >> 2);                               // inverse Gray code = digit
  const int ent = w == 0 ? 0 : (((w - 1) & ~1) ^ (((w - 1) & ~1) >> 1));   // entry point of sub-cube w
  const int dir = w == 0 ? 0 : (w & 1 ? trailing_ones(w) : trailing_ones(w - 1)) % 3;
  const int e2 = e ^ rotl(ent, d + 1), d2 = (d + dir + 1) % 3;
  return (uint32_t)(w | ((3 * e2 + d2) << 3));
}

// bits = bits per dimension; planar meshes (z extent 0) use a 2-D Morton key, the others a Hilbert key
// (IU_HILBERT = 0: Morton)
// idx = position of the query in the batch; bit 31 (PERM_HINT_BIT) is set when the query comes with a hint
// (i_cell >= 1), so that k_query reads the scattered 4-byte hint only for those: a scattered read costs a whole
// 128-byte DRAM line, which for an all-cold batch was a third of the kernel's DRAM reads (profiles/experiments/
// r02d_dram_by_kernel_form.txt).  hint == nullptr: no flag (nearest-centre calls, ray batches, tracer seeds).
constexpr uint32_t PERM_HINT_BIT = 0x80000000u;
__global__ void k_morton_keys(int64_t n, const double* __restrict__ r, const int32_t* __restrict__ hint, double lox, double loy,
                              double loz, double sx, double sy, double sz, int bits, int planar, uint32_t* __restrict__ keys,
                              uint32_t* __restrict__ idx) {
#if IU_HILBERT
  __shared__ uint8_t s_tab[192];
  if (!planar) {
    for (int t = threadIdx.x; t < 192; t += blockDim.x) s_tab[t] = (uint8_t)hilbert_table_entry(t >> 3, t & 7);
    __syncthreads();
  }
#endif
  const uint32_t maxv = (1u << bits) - 1u;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const uint32_t ix = quantise(__ldg(r + 3 * i), lox, sx, maxv);
    const uint32_t iy = quantise(__ldg(r + 3 * i + 1), loy, sy, maxv);
    uint32_t key;
    if (planar) key = spread2(ix) | (spread2(iy) << 1);
    else {
      const uint32_t iz = quantise(__ldg(r + 3 * i + 2), loz, sz, maxv);
      key = spread3(ix) | (spread3(iy) << 1) | (spread3(iz) << 2);
#if IU_HILBERT
      uint32_t h = 0, state = 0;
      for (int l = 3 * (bits - 1); l >= 0; l -= 3) {
        const uint32_t t = s_tab[state * 8 + ((key >> l) & 7u)];
        h = (h << 3) | (t & 7u);
        state = t >> 3;
      }
      key = h;
#endif
    }
    keys[i] = key;
    idx[i] = (uint32_t)i | (hint && LD_STREAM(hint + i) >= 1 ? PERM_HINT_BIT : 0u);
  }
}

// ------------------------------------------------------------------------------------------------
// device-side pieces of the hot path
// ------------------------------------------------------------------------------------------------
struct TreeView {
  const double* fat;     // per internal node (heap order) {split, left child's split, right child's split, bits(3 x 2-bit
                         // split dimension)}: one 32-byte load decides two levels of a descent
  const double* centers4;   // leaf blocks of 24 doubles: x[8], y[8], z[8] (padding slots = +inf)
  const int32_t* orig_of_dev;   // original id per slot: read only to break exact distance ties
  const float* box;      // 8 floats per node (internal and leaf): tight box of the centres below it
  int n_internal;
};

// Exact nearest centre (squared distance summed x, y, z without FMA, like kdtree2's leaf loop);
// ties -> lowest original cell index.  Depth-first traversal of the complete heap-indexed tree
// without a node stack:
//   * every descent step at level l stores a float LOWER bound (round-down) of the plane distance
//     (q[dim]-split)^2 in s_lb[l][thread] and marks level l in the `pending` bit trail iff the far
//     child may still hold a centre at distance <= best;
//   * after a leaf scan the deepest pending level whose bound still passes is popped, the walk
//     jumps straight to the far child there (sibling of the path node, from the heap index of the
//     current leaf by shifts) and descends again.  Pending bits always lie on the current
//     root-to-leaf path, so no per-level node ids are kept;
//   * the first descent and leaf scan run with the whole warp converged (uniform tree depth); the
//     first pop is preceded by a convergent filter of all levels against the first `best`;
//   * a popped far child is entered only if the tight box of the centres below it (box_lower_bound)
//     is within `best` as well;
//   * descents read "fat" nodes {split, both children's splits, dimensions}: one dependent 32-byte
//     load per two levels.
// By monotonicity of fp64 subtraction, multiplication and addition, (q[dim]-split)^2 is a lower
// bound of the COMPUTED squared distance of every centre behind the plane, and the float bound is
// <= it, so pruning with '>' never drops the true fp64 minimum or a tie.  Returns the 0-based
// device slot, -1 if none.
// q[k] for k = 0, 1, 2 without a branch (the compiler otherwise emits a divergent branch around
// the 64-bit moves in the innermost loop of the search)
__device__ __forceinline__ double select3(double qx, double qy, double qz, int k) {
#if IU_SELP
  double r;
  asm("{\n\t.reg .pred p0, p1;\n\tsetp.eq.s32 p0, %4, 0;\n\tsetp.eq.s32 p1, %4, 1;\n\t"
      "selp.f64 %0, %2, %3, p1;\n\tselp.f64 %0, %1, %0, p0;\n\t}"
      : "=&d"(r) : "d"(qx), "d"(qy), "d"(qz), "r"(k));
  return r;
#else
  return k == 0 ? qx : (k == 1 ? qy : qz);
#endif
}

struct NNState {
  double best;
  int besti;
};

__device__ __forceinline__ void scan_leaf(const TreeView& t, int node, double qx, double qy, double qz, NNState& st) {
  // two batches of four slots: x, y, z of four centres = three independent 32-byte loads.  The
  // common path is branch-free (min / select); exact ties only raise a flag and are sorted out below.
  const int base = (node - t.n_internal) * LEAF;
  if (IU_OOB(node < t.n_internal || node - t.n_internal >= g_lim[3])) return;
  const double* c = t.centers4 + 3 * (size_t)base;
  bool tie = false;
#pragma unroll
  for (int h = 0; h < LEAF; h += 4) {
    const D4 X = ld256(c + h), Y = ld256(c + LEAF + h), Z = ld256(c + 2 * LEAF + h);
    const double xs[4] = {X.x, X.y, X.z, X.w}, ys[4] = {Y.x, Y.y, Y.z, Y.w}, zs[4] = {Z.x, Z.y, Z.z, Z.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const double dx = xs[j] - qx, dy = ys[j] - qy, dz = zs[j] - qz;
      const double sd = (dx * dx + dy * dy) + dz * dz;
      tie |= sd == st.best;
      st.besti = sd < st.best ? base + h + j : st.besti;
      st.best = sd < st.best ? sd : st.best;
    }
  }
  if (tie) {
    // some slot equalled the running minimum (rare): among all slots of this leaf at the final
    // minimum and the incumbent, the lowest original cell id wins; padding slots (+inf, id 0) never do
    for (int j = 0; j < LEAF; ++j) {
      const double dx = __ldg(c + j) - qx, dy = __ldg(c + LEAF + j) - qy, dz = __ldg(c + 2 * LEAF + j) - qz;
      const double sd = (dx * dx + dy * dy) + dz * dz;
      if (sd == st.best && base + j != st.besti) {
        const int o = __ldg(t.orig_of_dev + base + j);
        if (o > 0 && (st.besti < 0 || o < __ldg(t.orig_of_dev + st.besti))) st.besti = base + j;
      }
    }
  }
}

// Float LOWER bound of the squared distance from q to the node's centre box, with directed rounding
// throughout: every centre c below the node has lo_k <= c_k <= hi_k (floats rounded outward), and
// with qd <= q <= qu (q rounded down / up to float) the per-dimension offset
// max(lo - qu, qd - hi, 0), each difference rounded down, is <= |c_k - q_k|; squares and sums are
// rounded down too.  The caller prunes when the bound exceeds `best` rounded UP to float: two distinct
// floats differ by 2^-24 relative, far more than the 2^-51 by which the fp64-computed distance of a
// centre can fall below the exact one, so no centre at a distance <= best (ties included) is dropped.
__device__ __forceinline__ float box_lower_bound(const float* b, double qx, double qy, double qz) {
  float lx, ly, lz, hx, hy, hz, p0, p1;
  asm volatile("ld.global.nc" IU_LTC_M ".v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(lx), "=f"(ly), "=f"(lz), "=f"(hx), "=f"(hy), "=f"(hz), "=f"(p0), "=f"(p1) : "l"(b));
  const float ox = fmaxf(fmaxf(__fsub_rd(lx, __double2float_ru(qx)), __fsub_rd(__double2float_rd(qx), hx)), 0.0f);
  const float oy = fmaxf(fmaxf(__fsub_rd(ly, __double2float_ru(qy)), __fsub_rd(__double2float_rd(qy), hy)), 0.0f);
  const float oz = fmaxf(fmaxf(__fsub_rd(lz, __double2float_ru(qz)), __fsub_rd(__double2float_rd(qz), hz)), 0.0f);
  return __fadd_rd(__fadd_rd(__fmul_rd(ox, ox), __fmul_rd(oy, oy)), __fmul_rd(oz, oz));
}

// Search state of one query between rounds (nn_begin, then nn_round until it returns false).
struct NNLane {
  NNState st;
  uint32_t pending;
  int node, lvl;
};

// first descent (best = +inf, every level stays pending), first leaf, then the filter of all levels
// against the first `best`: all of it runs with the lanes converged
__device__ __forceinline__ void nn_begin(const TreeView& t, float* lb, double qx,
                                         double qy, double qz, NNLane& L) {
  L.st = NNState{__longlong_as_double(0x7ff0000000000000LL), -1};
  L.pending = 0;
  int node = 0, lvl = 0;
  {
    auto step = [&](double s, int k) -> int {
      const double qk = select3(qx, qy, qz, k);
      const double d = qk - s;
      lb[lvl * QUERY_THREADS] = __double2float_rd(d * d);
      const int bit = qk < s ? 0 : 1;
      node = 2 * node + 1 + bit;
      ++lvl;
      return bit;
    };
    while (node < t.n_internal) {
      const D4 f = ld256(t.fat + 4 * (size_t)node);
      const int dims = __double2loint(f.w);
      const int b0 = step(f.x, dims & 3);
      if (node >= t.n_internal) break;
      step(b0 ? f.z : f.y, (dims >> (2 + 2 * b0)) & 3);
    }
  }
  scan_leaf(t, node, qx, qy, qz, L.st);
  const float bf = __double2float_ru(L.st.best);
  for (int l = 0; l < lvl; ++l)
    if (!(lb[l * QUERY_THREADS] > bf)) L.pending |= 1u << l;
  L.node = node; L.lvl = lvl;
}

// One backtracking round: pop the deepest pending level whose plane bound and far-child box still
// pass, descend from the far child to a leaf, scan it.  false = nothing left, the search is done.
__device__ __forceinline__ bool nn_round(const TreeView& t, float* lb, double qx,
                                         double qy, double qz, NNLane& L) {
  const float bf = __double2float_ru(L.st.best);
  int node = L.node, lvl = L.lvl;
  uint32_t pending = L.pending;
  bool found = false;
  while (pending) {
    const int l = 31 - __clz(pending);
    pending ^= 1u << l;
    if (!(lb[l * QUERY_THREADS] > bf)) {
      const int p1 = (node + 1) >> (lvl - (l + 1));   // 1-based heap index of the path node at level l+1
      const int far = (p1 ^ 1) - 1;                   // its sibling = the far child at level l
#if IU_NODE_BOX
      if (box_lower_bound(t.box + 8 * (size_t)far, qx, qy, qz) > bf) continue;
#endif
      node = far;
      lvl = l + 1;
      found = true;
      break;
    }
  }
  if (found) {
    {
      auto step = [&](double s, int k) -> int {
        const double qk = select3(qx, qy, qz, k);
        const double d = qk - s;
        const double d2 = d * d;
        lb[lvl * QUERY_THREADS] = __double2float_rd(d2);
        if (!(d2 > L.st.best)) pending |= 1u << lvl;
        const int bit = qk < s ? 0 : 1;
        node = 2 * node + 1 + bit;
        ++lvl;
        return bit;
      };
      while (node < t.n_internal) {
        const D4 f = ld256(t.fat + 4 * (size_t)node);
        const int dims = __double2loint(f.w);
        const int b0 = step(f.x, dims & 3);
        if (node >= t.n_internal) break;
        step(b0 ? f.z : f.y, (dims >> (2 + 2 * b0)) & 3);
      }
    }
    scan_leaf(t, node, qx, qy, qz, L.st);
  }
  L.node = node; L.lvl = lvl; L.pending = pending;
  return found;
}

__device__ __forceinline__ int nearest_center(const TreeView& t, float* s_lb,
                                              double qx, double qy, double qz) {
  float* lb = s_lb + threadIdx.x;   // lb[l * QUERY_THREADS]
  NNLane L;
  nn_begin(t, lb, qx, qy, qz, L);
  while (nn_round(t, lb, qx, qy, qz, L)) {}
  return L.st.besti;
}

// ------------------------------------------------------------------------------------------------
// Voxel table of the nearest-centre search (device.h: VoxelGrid).
//
// k_build_voxels, one warp per voxel B (the voxel's box inflated by VOX_DELTA voxel sizes per side, which covers the
// rounding of the query's voxel index): gathers with a depth-first traversal of the kd-tree every centre c with
//     mindist^2(c, B) <= (1 + eps) * D2,   D2 = min over centres c' of maxdist^2(c', B)
// (for q in B the nearest centre is at most maxdist(c', B) away), then drops every c for which some gathered c' has
//     min over q in B of (|q - c|^2 - |q - c'|^2)  >  eps * maxdist^2(c, B)
// (the difference is linear in q, so its minimum over the box sits at the corner picked per dimension by the sign of
// c - c').  A dropped centre is, for every q of the voxel, farther than some other centre by a relative margin
// eps = 1e-9, seven orders of magnitude above the rounding error of the fp64 distance the search computes, so the list
// keeps every centre that can be nearest OR tie for nearest in computed arithmetic.  Voxels with more than VOX_K
// survivors (or more than VOX_GATHER gathered) are flagged VOX_TREE.
// ------------------------------------------------------------------------------------------------
constexpr int VOX_GATHER = 160;
constexpr double VOX_EPS = 1e-9, VOX_DELTA = 1e-3;

__device__ __forceinline__ double warp_min(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmin(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// One warp per BLOCK of 2 x 2 x 2 voxels.  The tree is traversed once for the union U of the block's (inflated) voxel
// boxes - all 32 lanes of a warp evaluate the same box tests, so a traversal per voxel spent 41 % of the kernel's
// instructions on redundant work (profiles/r02w_k_build_voxels_metrics.txt) - and the gathered centres are then filtered
// for each voxel B of the block.  This is the same list a traversal for B alone produces: with
// D2(B) = min_c maxdist^2(c, B) <= D2(U), every centre within (1 + eps) D2(B) of B is within (1 + eps) D2(U) of U
// (mindist(c, U) <= mindist(c, B)), and so is the centre that attains D2(B); a gathered centre outside that set can never
// remove one inside it (it would have to be closer than it for every point of B).  A block whose union gathers more than
// VOX_GATHER centres falls back to one traversal per voxel.
__global__ void __launch_bounds__(256) k_build_voxels(TreeView t, VoxelGrid vg, int64_t n_vox, uint32_t* __restrict__ lists,
                                                      unsigned long long* __restrict__ stats) {
  __shared__ double s_c[8][3][VOX_GATHER];
  __shared__ int s_slot[8][VOX_GATHER];
  __shared__ int s_stack[8][72];
  __shared__ double s_sd2[8][72];
  __shared__ unsigned char s_alive[8][VOX_GATHER];
  __shared__ int s_out[8][VOX_K];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const double inf = __longlong_as_double(0x7ff0000000000000LL);
  unsigned long long n_flagged = 0, n_entries = 0;
  (void)n_vox;
  int nb[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) nb[k] = (vg.n[k] + 1) / 2;
  const int64_t n_blocks = (int64_t)nb[0] * nb[1] * nb[2];

  double D2 = inf;
  int count = 0;
  bool overflow = false;

  // gathers into s_c / s_slot every centre within (1 + eps) of the running D2 of the box [blo, bhi] (centre bc)
  auto gather = [&](const double* blo, const double* bhi, const double* bc) {
    D2 = inf; count = 0; overflow = false;
    int sp = 0;
    // A node is pushed together with the squared distance from the box to the box of its centres, computed when its
    // PARENT is expanded: the parent's fat node and both children's boxes (64 contiguous bytes) are fetched in one round
    // trip, and children beyond the current bound are never pushed.  D2 only shrinks, so the stored distance is
    // re-checked against the bound of the moment when the node is popped.
    auto box_d2 = [&](const float* b) {
      float v[8];
      asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]) : "l"(b));
      double d2 = 0;
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const double g = fmax(fmax((double)v[k] - bhi[k], blo[k] - (double)v[3 + k]), 0.0);
        d2 += g * g;
      }
      return d2;
    };
    if (lane == 0) { s_stack[w][0] = 0; s_sd2[w][0] = 0.0; }
    sp = 1;
    __syncwarp();
    while (sp > 0) {
      --sp;
      const int node = s_stack[w][sp];
      const double d2 = s_sd2[w][sp];
      __syncwarp();
      if (d2 > D2 * (1.0 + VOX_EPS)) continue;
      if (node >= t.n_internal) {
        const int base = (node - t.n_internal) * LEAF;
        bool keep = false;
        double cx = 0, cy = 0, cz = 0, maxd2 = inf;
        if (lane < LEAF) {
          const double* c = t.centers4 + 3 * (size_t)base + lane;
          cx = __ldg(c); cy = __ldg(c + LEAF); cz = __ldg(c + 2 * LEAF);
          if (cx < inf) {
            const double ax = fmax(fabs(cx - blo[0]), fabs(cx - bhi[0])), ay = fmax(fabs(cy - blo[1]), fabs(cy - bhi[1])),
                         az = fmax(fabs(cz - blo[2]), fabs(cz - bhi[2]));
            maxd2 = (ax * ax + ay * ay) + az * az;
            keep = true;
          }
        }
        D2 = fmin(D2, warp_min(maxd2));
        if (keep) {
          const double gx = fmax(fmax(blo[0] - cx, cx - bhi[0]), 0.0), gy = fmax(fmax(blo[1] - cy, cy - bhi[1]), 0.0),
                       gz = fmax(fmax(blo[2] - cz, cz - bhi[2]), 0.0);
          keep = (gx * gx + gy * gy) + gz * gz <= D2 * (1.0 + VOX_EPS);
        }
        const unsigned m = __ballot_sync(0xffffffffu, keep);
        const int pos = count + __popc(m & ((1u << lane) - 1u));
        if (keep && pos < VOX_GATHER) { s_c[w][0][pos] = cx; s_c[w][1][pos] = cy; s_c[w][2][pos] = cz; s_slot[w][pos] = base + lane; }
        count += __popc(m);
        if (count > VOX_GATHER) { overflow = true; count = VOX_GATHER; }
      } else {
        const D4 f = ld256(t.fat + 4 * (size_t)node);
        const double dl = box_d2(t.box + 8 * (size_t)(2 * node + 1)), dr = box_d2(t.box + 8 * (size_t)(2 * node + 2));
        const int dim = __double2loint(f.w) & 3;
        const int near = (dim == 0 ? bc[0] : dim == 1 ? bc[1] : bc[2]) < f.x ? 0 : 1;
        const double bound = D2 * (1.0 + VOX_EPS);
        const int c_far = 2 * node + 2 - near, c_near = 2 * node + 1 + near;
        const double d_far = near ? dl : dr, d_near = near ? dr : dl;
        if (!(d_far > bound)) { if (lane == 0) { s_stack[w][sp] = c_far; s_sd2[w][sp] = d_far; } ++sp; }
        if (!(d_near > bound)) { if (lane == 0) { s_stack[w][sp] = c_near; s_sd2[w][sp] = d_near; } ++sp; }
        __syncwarp();
      }
    }
    __syncwarp();
  };

  // the list of voxel `vox` = box [blo, bhi] from the gathered centres: within (1 + eps) D2v of the box and not beaten by
  // another gathered centre over the whole box
  auto emit = [&](int64_t vox, const double* blo, const double* bhi, double D2v, bool ovf) {
    int n_alive = 0;
    if (!ovf) {
      for (int i = lane; i < count; i += 32) {
        const double cx = s_c[w][0][i], cy = s_c[w][1][i], cz = s_c[w][2][i];
        const double gx = fmax(fmax(blo[0] - cx, cx - bhi[0]), 0.0), gy = fmax(fmax(blo[1] - cy, cy - bhi[1]), 0.0),
                     gz = fmax(fmax(blo[2] - cz, cz - bhi[2]), 0.0);
        bool alive = (gx * gx + gy * gy) + gz * gz <= D2v * (1.0 + VOX_EPS);
        const double ax = fmax(fabs(cx - blo[0]), fabs(cx - bhi[0])), ay = fmax(fabs(cy - blo[1]), fabs(cy - bhi[1])),
                     az = fmax(fabs(cz - blo[2]), fabs(cz - bhi[2]));
        const double margin = VOX_EPS * ((ax * ax + ay * ay) + az * az);
        for (int j = 0; j < count && alive; ++j) {
          const double ox = s_c[w][0][j], oy = s_c[w][1][j], oz = s_c[w][2][j];
          const double qx = cx - ox > 0 ? bhi[0] : blo[0], qy = cy - oy > 0 ? bhi[1] : blo[1], qz = cz - oz > 0 ? bhi[2] : blo[2];
          const double fi = ((qx - cx) * (qx - cx) + (qy - cy) * (qy - cy)) + (qz - cz) * (qz - cz);
          const double fj = ((qx - ox) * (qx - ox) + (qy - oy) * (qy - oy)) + (qz - oz) * (qz - oz);
          if (fi - fj > margin) alive = false;
        }
        s_alive[w][i] = alive ? 1 : 0;
      }
      __syncwarp();
      for (int i0 = 0; i0 < count; i0 += 32) {   // survivors in traversal order (nearest leaves first)
        const int i = i0 + lane;
        const bool a = i < count && s_alive[w][i];
        const unsigned m = __ballot_sync(0xffffffffu, a);
        const int pos = n_alive + __popc(m & ((1u << lane) - 1u));
        if (a && pos < VOX_K) s_out[w][pos] = s_slot[w][i];
        n_alive += __popc(m);
      }
      __syncwarp();
    }
    const bool flagged = ovf || n_alive > VOX_K || n_alive == 0;
    if (lane < VOX_K) {
      uint32_t e;
      if (flagged) e = lane == 0 ? VOX_TREE : VOX_END;
      else if (lane < n_alive) e = (uint32_t)s_out[w][lane];
      else if (lane < ((n_alive + 3) & ~3)) e = (uint32_t)s_out[w][n_alive - 1];   // pad the last group with a repeat
      else e = VOX_END;
      lists[(size_t)vox * VOX_K + lane] = e;
    }
    if (flagged) ++n_flagged; else n_entries += (unsigned long long)n_alive;
    __syncwarp();
  };

  // inflated box of the voxel range [i0, i1) per dimension (VOX_DELTA voxel sizes per side cover the rounding of the
  // query's voxel index)
  auto range_box = [&](const int* i0, const int* i1, double* blo, double* bhi, double* bc) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      if (vg.flat_mask >> k & 1) { blo[k] = bhi[k] = bc[k] = vg.flat_val[k]; }
      else {
        blo[k] = vg.lo[k] + ((double)i0[k] - VOX_DELTA) * vg.h[k];
        bhi[k] = vg.lo[k] + ((double)i1[k] + VOX_DELTA) * vg.h[k];
        bc[k] = 0.5 * (blo[k] + bhi[k]);
      }
    }
  };

  for (int64_t blk = (int64_t)blockIdx.x * 8 + w; blk < n_blocks; blk += (int64_t)gridDim.x * 8) {
    int b0[3], b1[3];
    b0[0] = 2 * (int)(blk % nb[0]); b0[1] = 2 * (int)((blk / nb[0]) % nb[1]); b0[2] = 2 * (int)(blk / ((int64_t)nb[0] * nb[1]));
#pragma unroll
    for (int k = 0; k < 3; ++k) b1[k] = min(b0[k] + 2, vg.n[k]);
    double ulo[3], uhi[3], uc[3];
    range_box(b0, b1, ulo, uhi, uc);
    gather(ulo, uhi, uc);
    const bool block_overflow = overflow;
    for (int vz = b0[2]; vz < b1[2]; ++vz)
      for (int vy = b0[1]; vy < b1[1]; ++vy)
        for (int vx = b0[0]; vx < b1[0]; ++vx) {
          const int v0[3] = {vx, vy, vz}, v1[3] = {vx + 1, vy + 1, vz + 1};
          double blo[3], bhi[3], bc[3];
          range_box(v0, v1, blo, bhi, bc);
          const int64_t vox = ((int64_t)vz * vg.n[1] + vy) * vg.n[0] + vx;
          if (block_overflow) {          // too many centres around the block: one traversal for this voxel alone
            gather(blo, bhi, bc);
            emit(vox, blo, bhi, D2, overflow);
            continue;
          }
          double m = inf;                // D2 of this voxel from the gathered centres (the centre that attains it is among them)
          for (int i = lane; i < count; i += 32) {
            const double cx = s_c[w][0][i], cy = s_c[w][1][i], cz = s_c[w][2][i];
            const double ax = fmax(fabs(cx - blo[0]), fabs(cx - bhi[0])), ay = fmax(fabs(cy - blo[1]), fabs(cy - bhi[1])),
                         az = fmax(fabs(cz - blo[2]), fabs(cz - bhi[2]));
            m = fmin(m, (ax * ax + ay * ay) + az * az);
          }
          emit(vox, blo, bhi, warp_min(m), false);
        }
  }
  if (lane == 0) {
    if (n_flagged) atomicAdd(stats, n_flagged);
    if (n_entries) atomicAdd(stats + 1, n_entries);
  }
}

// Nearest centre through the voxel table: the exact fp64 distance of scan_leaf for every listed candidate, ties to the
// lowest original id.  Returns the 0-based device slot, or -2 when the query must search the tree (outside the grid,
// NaN, off-plane query of a planar mesh, flagged voxel).
__device__ __forceinline__ int nn_voxel(const VoxelGrid& V, const double* __restrict__ cen32, double qx, double qy, double qz) {
  const double tx = (qx - V.lo[0]) * V.inv[0], ty = (qy - V.lo[1]) * V.inv[1], tz = (qz - V.lo[2]) * V.inv[2];
  bool ok = tx >= 0.0 && tx < (double)V.n[0] && ty >= 0.0 && ty < (double)V.n[1] && tz >= 0.0 && tz < (double)V.n[2];
  if (V.flat_mask) {
    if ((V.flat_mask & 1) && qx != V.flat_val[0]) ok = false;
    if ((V.flat_mask & 2) && qy != V.flat_val[1]) ok = false;
    if ((V.flat_mask & 4) && qz != V.flat_val[2]) ok = false;
  }
  if (!ok) return -2;
  const size_t vox = ((size_t)(int)tz * V.n[1] + (size_t)(int)ty) * V.n[0] + (size_t)(int)tx;
  if (IU_OOB((long long)vox >= g_lim[2])) return -2;
  const uint4* L = reinterpret_cast<const uint4*>(V.lists) + vox * (VOX_K / 4);
  double best = __longlong_as_double(0x7ff0000000000000LL);
  int besti = -2, besto = 0x7fffffff;
#pragma unroll 1
  for (int g = 0; g < VOX_K / 4; ++g) {
    const uint4 e = ld128u(L + g);
    if (e.x >= VOX_TREE) break;   // VOX_END: list complete; VOX_TREE (first group only): besti stays -2
    const uint32_t sl[4] = {e.x, e.y, e.z, e.w};
    if (IU_OOB((long long)e.x >= g_lim[0] || (long long)e.y >= g_lim[0] || (long long)e.z >= g_lim[0] || (long long)e.w >= g_lim[0])) break;
    D4 c[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) c[j] = ld256(cen32 + 4 * (size_t)sl[j]);   // (predicating off the padded repeats: fewer wavefronts, +7 % instructions, slower)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const double dx = c[j].x - qx, dy = c[j].y - qy, dz = c[j].z - qz;
      const double sd = (dx * dx + dy * dy) + dz * dz;
      const int o = __double2loint(c[j].w);
      const bool take = sd < best || (sd == best && o < besto);
      best = take ? sd : best;
      besti = take ? (int)sl[j] : besti;
      besto = take ? o : besto;
    }
  }
  return besti;
}

// iu_get_cell_through_neighbors (:667-725) + get_cell_intersection (:729-764).
// Cells are 1-based DEVICE ids.  mask = icell_data column (original cell order) or null.
template <int PPC>
__device__ __forceinline__ void walk_ray(const double* __restrict__ rec, V3 r0, V3 r1, int ic0, int max_steps,
                                         const int32_t* __restrict__ mask, int& ic1, V3& r_p, int& n_steps, int& status,
                                         bool& early) {
  using L = RecLayout<PPC>;
  const V3 dvec = r1 - r0;
  double distance_left = norm2_3(dvec);
  status = 0;
  early = false;
  ic1 = ic0;
  if (distance_left < TINY_DISTANCE) { early = true; return; }   // :689-692
  const V3 u{dvec.x / distance_left, dvec.y / distance_left, dvec.z / distance_left};   // :694
  r_p = r0;
  int mask0 = 0;
  if (mask) mask0 = __ldg(mask + (__ldg(reinterpret_cast<const int32_t*>(rec + (size_t)(ic0 - 1) * L::STRIDE) + L::ORG_I) - 1));
  for (n_steps = 1; n_steps <= max_steps; ++n_steps) {
    if (IU_OOB(ic1 < 1 || ic1 > g_lim[0])) { status = -1; ic1 = 0; break; }
    const double* R = rec + (size_t)(ic1 - 1) * L::STRIDE;
    double face_distance = HUGE_DISTANCE;
    int i_face = -1;
    // All faces are evaluated without branching (a division whose face fails `pdn > 0` is simply
    // discarded): the same operations as :744-760 for the faces that count, and no divergence
    // between lanes whose rays leave through different faces.
    if (PPC == 4 && IU_WALK_256) {   // 96 + 96 bytes as six 32-byte loads
      V3 fn[4], rf[4];
      ld3x4(R + L::NRM, fn);
      ld3x4(R + L::PTS, rf);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const double pdn = dot3(u, fn[k]);
        const double dist = dot3(rf[k] - r_p, fn[k]) / pdn;
        if (pdn > 0 && dist < face_distance) { face_distance = dist; i_face = k; }
      }
    } else if (PPC == 4) {           // faces two at a time: 6 + 6 doubles as 16-byte loads
#pragma unroll
      for (int h = 0; h < 4; h += 2) {
        V3 fn[2], rf[2];
        ld3n<2>(R + L::NRM + 3 * h, fn);
        ld3n<2>(R + L::PTS + 3 * h, rf);
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const double pdn = dot3(u, fn[j]);
          const double dist = dot3(rf[j] - r_p, fn[j]) / pdn;
          if (pdn > 0 && dist < face_distance) { face_distance = dist; i_face = h + j; }
        }
      }
    } else {
      V3 fn[3], rf[3];
      ld3n<3>(R + L::NRM, fn);
      ld3n<3>(R + L::PTS, rf);
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const double pdn = dot3(u, fn[k]);
        const double dist = dot3(rf[k] - r_p, fn[k]) / pdn;
        if (pdn > 0 && dist < face_distance) { face_distance = dist; i_face = k; }
      }
    }
    r_p = {r_p.x + face_distance * u.x, r_p.y + face_distance * u.y, r_p.z + face_distance * u.z};   // :762
    distance_left = distance_left - face_distance;
    if (distance_left > 0) {
      ic1 = __ldg(reinterpret_cast<const int32_t*>(R) + L::NBR_I + i_face);
      if (ic1 < 1) { status = -1; break; }
      if (mask) {
        const int o1 = __ldg(reinterpret_cast<const int32_t*>(rec + (size_t)(ic1 - 1) * L::STRIDE) + L::ORG_I);
        if (mask0 != __ldg(mask + (o1 - 1))) { status = 1; break; }
      }
    } else {
      break;
    }
  }
}

// interpolate_triangle (:529-551), interpolate_quad (:591-641), interpolate_tetrahedron (:555-586):
// computes the per-vertex weights once, then applies them to every requested variable.
template <int CT> struct Weights;
template <> struct Weights<IU_TETRA_CT> {
  double w[4], inv;
  __device__ __forceinline__ void setup(const V3* p, double vol, V3 r) {
    const V3 v1r = r - p[0], v2r = r - p[1];
    const V3 v12 = p[1] - p[0], v13 = p[2] - p[0], v14 = p[3] - p[0], v23 = p[2] - p[1], v24 = p[3] - p[1];
    w[0] = stp(v2r, v24, v23);
    w[1] = stp(v1r, v13, v14);
    w[2] = stp(v1r, v14, v12);
    w[3] = stp(v1r, v12, v13);
    inv = 1 / (6 * vol);
  }
  __device__ __forceinline__ double apply(const double* d) const {
    return (((d[0] * w[0] + d[1] * w[1]) + d[2] * w[2]) + d[3] * w[3]) * inv;
  }
};
template <> struct Weights<IU_TRIANGLE_CT> {
  double w[3], inv;
  __device__ __forceinline__ void setup(const V3* p, double area, V3 r) {
    w[0] = 0.5 * norm2_3(cross(r - p[1], r - p[2]));
    w[1] = 0.5 * norm2_3(cross(r - p[2], r - p[0]));
    w[2] = 0.5 * norm2_3(cross(r - p[0], r - p[1]));
    inv = 1 / area;
  }
  __device__ __forceinline__ double apply(const double* d) const {
    return ((d[0] * w[0] + d[1] * w[1]) + d[2] * w[2]) * inv;
  }
};
template <> struct Weights<IU_QUAD_CT> {
  double c1, c2;
  __device__ __forceinline__ void setup(const V3* p, double, V3 r) {
    const double tiny_value = 1e-20;
    const V3 q = r - p[0];
    const V3 b1 = p[1] - p[0];
    const V3 b2 = p[3] - p[0];
    const V3 b3 = ((p[0] - p[1]) - p[3]) + p[2];
    const double A = cross_z(b2, b3);
    const double B = cross_z(b3, q) - cross_z(b1, b2);
    const double C = cross_z(b1, q);
    const double discrim = B * B - 4 * A * C;
    if (fabs(A) < tiny_value) c2 = -C / B;
    else c2 = 0.5 * (-B - sqrt(discrim)) / A;
    const double dx = b1.x + c2 * b3.x, dy = b1.y + c2 * b3.y, dz = b1.z + c2 * b3.z;
    // maxloc(abs(denom)): first maximum
    int dim = 0; double dm = fabs(dx);
    if (fabs(dy) > dm) { dim = 1; dm = fabs(dy); }
    if (fabs(dz) > dm) { dim = 2; }
    const double qd = dim == 0 ? q.x : (dim == 1 ? q.y : q.z);
    const double b2d = dim == 0 ? b2.x : (dim == 1 ? b2.y : b2.z);
    const double dd = dim == 0 ? dx : (dim == 1 ? dy : dz);
    c1 = (qd - b2d * c2) / dd;
  }
  __device__ __forceinline__ double apply(const double* d) const {
    const double t1 = d[0] * (1 - c1) + d[1] * c1;
    const double t2 = d[3] * (1 - c1) + d[2] * c1;
    return t1 * (1 - c2) + t2 * c2;
  }
};

struct QueryKernelParams {
  TreeView tree;
  VoxelGrid vox;
  const double* cen32;
  const double* rec;
  const int32_t* dev_of_orig;
  const int32_t* orig_of_dev;
  const double* pd;
  const double* cd;
  const int32_t* icd;
  unsigned long long* counters;
  const double* r;
  const uint32_t* perm;
  int32_t* i_cell;
  double* out_f64;
  double* out2_f64;               // QM_INTERP_CELL: the cell-data scalar of each query
  int32_t* out_i32;
  int64_t n;
  int n_cells;
  int cell_var;                   // QM_INTERP_CELL: 0-based cell_data column
  int pd_stride;
  int n_vars;
  int max_steps;
  int pairs;                      // 1 / 2: the requested variables come in adjacent, aligned groups of 2 / 4 columns
  unsigned long long* ticket;     // work counter of the dynamically scheduled variants (zeroed before the launch)
  int sched;
  int tiles_per_warp;             // SCHED = 1: a warp retires after this many tiles (0: never = persistent CTAs)
  int ivars[MAX_VARS];
};

// the kd-tree search behind one more function boundary for the lane-refill kernel (marking it __noinline__ crashes
// nvcc 12.9's front end on this file, so it is an ordinary inlinable function)
__device__ int nearest_center_call(const double* fat, const double* centers4, const int32_t* orig_of_dev,
                                                const float* box, int n_internal, float* s_lb, double qx, double qy, double qz) {
  const TreeView t{fat, centers4, orig_of_dev, box, n_internal};
  return nearest_center(t, s_lb, qx, qy, qz);
}

// find_nearby_cell_kdtree (:272-288): the voxel table where it applies, else the kd-tree search
__device__ __forceinline__ int nearest_slot(const QueryKernelParams& P, float* s_lb, double qx, double qy, double qz) {
  int slot = -2;
  if (P.vox.lists) slot = nn_voxel(P.vox, P.cen32, qx, qy, qz);
  if (slot == -2) {
    if (P.vox.lists) atomicAdd(P.counters + CNT_NN_TREE, 1ULL);
    slot = nearest_center(P.tree, s_lb, qx, qy, qz);
  }
  return slot;
}
__device__ __forceinline__ int nearest_slot_call(const QueryKernelParams& P, float* s_lb, double qx, double qy, double qz) {
  int slot = -2;
  if (P.vox.lists) slot = nn_voxel(P.vox, P.cen32, qx, qy, qz);
  if (slot == -2) {
    if (P.vox.lists) atomicAdd(P.counters + CNT_NN_TREE, 1ULL);
    slot = nearest_center_call(P.tree.fat, P.tree.centers4, P.tree.orig_of_dev, P.tree.box, P.tree.n_internal, s_lb, qx, qy, qz);
  }
  return slot;
}

// Everything after the start cell is known: ray walk from the start cell's centre (iu_get_cell,
// :429-432), then the mode's output.  `start` is a 1-based device id; for QM_INTERP_ICELL it is the
// containing cell itself (no walk).
template <int CT, int MODE, int NV>
__device__ __forceinline__ void emit_result(const QueryKernelParams& P, int64_t i, double qx, double qy, double qz, int ic);

template <int CT, int MODE, int NV>
__device__ __forceinline__ void finish_query(const QueryKernelParams& P, int64_t i, double qx, double qy, double qz, int start) {
  constexpr int PPC = CT == IU_TRIANGLE_CT ? 3 : 4;
  int ic = start;
  if (IU_OOB(start < 1 || start > g_lim[0])) return;
  if (MODE != QM_INTERP_ICELL) {
    const D4 c0 = ld256(P.cen32 + 4 * (size_t)(start - 1));
    const V3 r0{c0.x, c0.y, c0.z};                                                    // :429
    V3 r_p; int n_steps = 0, status; bool early;
    walk_ray<PPC>(P.rec, r0, V3{qx, qy, qz}, start, P.max_steps, nullptr, ic, r_p, n_steps, status, early);
    if (!early && n_steps > P.max_steps) {
      ic = IU_WALK_LIMIT;
      atomicAdd(P.counters + CNT_WALK_LIMIT, 1ULL);
    }
  }
  emit_result<CT, MODE, NV>(P, i, qx, qy, qz, ic);
}

// The walk has ended in device cell ic (1-based; < 1: left the domain / step cap): the mode's outputs.
template <int CT, int MODE, int NV>
__device__ __forceinline__ void emit_result(const QueryKernelParams& P, int64_t i, double qx, double qy, double qz, int ic) {
  constexpr int PPC = CT == IU_TRIANGLE_CT ? 3 : 4;
  using L = RecLayout<PPC>;
  if (MODE != QM_INTERP_ICELL && ic < 1) { ST_STREAM(P.i_cell + i, ic); atomicAdd(P.counters + CNT_NOT_FOUND, 1ULL); return; }   // :492
  const double* R = P.rec + (size_t)(ic - 1) * L::STRIDE;
  const int32_t* RI = reinterpret_cast<const int32_t*>(R);
  int orig, pid[4];
  double vol;
  if constexpr (PPC == 4) {   // vertex ids, volume and original id share one 32-byte sector
    const D4 tail = ld256(R + L::TAIL);
    pid[0] = __double2loint(tail.x); pid[1] = __double2hiint(tail.x);
    pid[2] = __double2loint(tail.y); pid[3] = __double2hiint(tail.y);
    vol = tail.z;
    orig = __double2loint(tail.w);
  } else {
    orig = __ldg(RI + L::ORG_I);
    const int4 c = __ldg(reinterpret_cast<const int4*>(RI + L::CEL_I));
    pid[0] = c.x; pid[1] = c.y; pid[2] = c.z; pid[3] = c.w;
    vol = __ldg(R + L::VOL);
  }
  if (IU_OOB(orig < 1 || orig > P.n_cells || pid[0] < 0 || pid[0] >= g_lim[1] || pid[1] < 0 || pid[1] >= g_lim[1] ||
             pid[2] < 0 || pid[2] >= g_lim[1] || (PPC == 4 && (pid[3] < 0 || pid[3] >= g_lim[1])))) return;
  if (MODE != QM_INTERP_ICELL && MODE != QM_INTERP_CELL) ST_STREAM(P.i_cell + i, orig);
  if (MODE == QM_CELL_F64) {
    ST_STREAM(P.out_f64 + i, __ldg(P.cd + (size_t)P.ivars[0] * P.n_cells + (orig - 1)));          // :447
  } else if (MODE == QM_CELL_I32) {
    ST_STREAM(P.out_i32 + i, __ldg(P.icd + (size_t)P.ivars[0] * P.n_cells + (orig - 1)));         // :461
  } else if (MODE == QM_INTERP || MODE == QM_INTERP_ICELL || MODE == QM_INTERP_CELL) {   // :498-527
    V3 p[PPC];
    if (PPC == 4) ld3x4(R + L::PTS, p); else ld3n<PPC>(R + L::PTS, p);
    Weights<CT> W;
    W.setup(p, vol, V3{qx, qy, qz});
    double* out = P.out_f64 + (size_t)i * P.n_vars;
    if (NV > 0) {
      double res[NV > 0 ? NV : 1];
      if (NV % 4 == 0 && P.pairs == 2) {   // aligned groups of four adjacent columns: one 32-byte load per vertex
#pragma unroll
        for (int v = 0; v < NV; v += 4) {
          double d0[4], d1[4], d2[4], d3[4];
#pragma unroll
          for (int k = 0; k < PPC; ++k) {
            const D4 x = ld256(P.pd + (size_t)pid[k] * P.pd_stride + P.ivars[v]);
            d0[k] = x.x; d1[k] = x.y; d2[k] = x.z; d3[k] = x.w;
          }
          res[v] = W.apply(d0); res[v + 1] = W.apply(d1); res[v + 2] = W.apply(d2); res[v + 3] = W.apply(d3);
        }
      } else if (NV % 2 == 0 && P.pairs) {   // adjacent column pairs: one 16-byte load per vertex and pair
#pragma unroll
        for (int v = 0; v < NV; v += 2) {
          double d0[4], d1[4];
#pragma unroll
          for (int k = 0; k < PPC; ++k) {
            const double2 x = __ldg(reinterpret_cast<const double2*>(P.pd + (size_t)pid[k] * P.pd_stride + P.ivars[v]));
            d0[k] = x.x; d1[k] = x.y;
          }
          res[v] = W.apply(d0);
          res[v + 1] = W.apply(d1);
        }
      } else {
#pragma unroll
        for (int v = 0; v < NV; ++v) {
          double d[4];
#pragma unroll
          for (int k = 0; k < PPC; ++k) d[k] = __ldg(P.pd + (size_t)pid[k] * P.pd_stride + P.ivars[v]);
          res[v] = W.apply(d);
        }
      }
      if (NV % 4 == 0 && (reinterpret_cast<uintptr_t>(out) & 31) == 0) {   // one 32-byte streaming store per four values
#pragma unroll
        for (int v = 0; v < NV; v += 4)
          asm volatile("st.global.cs.v4.f64 [%0], {%1,%2,%3,%4};" :: "l"(out + v), "d"(res[v]), "d"(res[v + 1]), "d"(res[v + 2]), "d"(res[v + 3]) : "memory");
      } else if (NV % 2 == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
#pragma unroll
        for (int v = 0; v < NV; v += 2) ST_STREAM(reinterpret_cast<double2*>(out + v), make_double2(res[v], res[v + 1]));
      } else {
#pragma unroll
        for (int v = 0; v < NV; ++v) ST_STREAM(out + v, res[v]);
      }
    } else {
      for (int v = 0; v < P.n_vars; ++v) {
        double d[4];
#pragma unroll
        for (int k = 0; k < PPC; ++k) d[k] = __ldg(P.pd + (size_t)pid[k] * P.pd_stride + P.ivars[v]);
        ST_STREAM(out + v, W.apply(d));
      }
    }
    if (MODE == QM_INTERP_CELL) {
      // The second call of the composition, iu_get_cell_scalar_at with the cell just found as the hint (:437-448):
      // iu_get_cell walks again, from that cell's centre (:424-432).  The record is still in L1 and the walk ends in
      // its first step, except for a query within rounding of a face, which may end in the neighbour - exactly as the
      // two-call sequence would.
      const D4 c1 = ld256(P.cen32 + 4 * (size_t)(ic - 1));
      V3 r_p; int ic2, n_steps = 0, status; bool early;
      walk_ray<PPC>(P.rec, V3{c1.x, c1.y, c1.z}, V3{qx, qy, qz}, ic, P.max_steps, nullptr, ic2, r_p, n_steps, status, early);
      if (!early && n_steps > P.max_steps) { ic2 = IU_WALK_LIMIT; atomicAdd(P.counters + CNT_WALK_LIMIT, 1ULL); }
      if (ic2 < 1) { ST_STREAM(P.i_cell + i, ic2); atomicAdd(P.counters + CNT_NOT_FOUND, 1ULL); return; }   // :446, res untouched
      const int orig2 = ic2 == ic ? orig : __ldg(reinterpret_cast<const int32_t*>(P.rec + (size_t)(ic2 - 1) * L::STRIDE) + L::ORG_I);
      ST_STREAM(P.i_cell + i, orig2);
      ST_STREAM(P.out2_f64 + i, __ldg(P.cd + (size_t)P.cell_var * P.n_cells + (orig2 - 1)));              // :447
    }
  }
}

// THE hot kernel.  Persistent CTAs (grid = SMs x resident CTAs); one query per thread, processed in curve order through
// `perm`; the dynamic shared memory holds the per-thread plane bounds of the kd-tree search (nearest_center), which only
// the queries without a voxel list run.
// SCHED = 0: static grid-stride assignment (unsorted batches).  SCHED = 1 (sorted batches): every
// warp draws tiles of IU_STILE (64) consecutive sorted queries from a global ticket counter, so that all
// queries in flight on the GPU stay within one narrow window of the Morton order whatever the
// speed differences between warps and SMs.
// A walk whose FIRST iteration reads precomputed numerators (p_k - centre) . n_k instead of the start cell's vertices
// (8 loads from one line instead of 12 from two, bit-identical) was built and dropped: the vertices are needed by the
// interpolation anyway and 77 % of the queries end in their start cell, so the loads only moved (7.76 vs 7.57 ms,
// profiles/experiments/r02j_first_step_precomputed.txt).
// Designs that were built, measured on B200 and dropped (profiles/README.md): a three-phase
// tiled kernel with a lane queue for the backtracking phase (fewer instructions, but the lost
// intra-warp locality costs more L1 wavefronts than it saves), and a warp-packet kd-tree search
// (uniform 32-byte loads still pay 8 LSU data-pipe wavefronts each for register write-back); and a
// lane-refill scheduler (finished lanes take the next queries of the tile once 16 are idle, one
// backtracking round per iteration): lane utilisation did not rise (14.0 vs 14.7 of 32) because the
// spread is inside a round (descent lengths, pop loops), not in the number of rounds: 18.9 vs 15.4 ms.
template <int CT, int MODE, int NV, int SCHED>
__global__ void __launch_bounds__(QUERY_THREADS, IU_QUERY_MIN_BLOCKS) k_query(const QueryKernelParams P) {
  extern __shared__ float s_lb[];   // [tree depth][QUERY_THREADS] plane-distance lower bounds (nearest_center)
  // SCHED = 1: warps draw STILE consecutive (sorted) queries at a time from the ticket counter
#ifndef IU_STILE
#define IU_STILE 64    // queries per ticket tile (with 3 CTAs per SM 64 beats 128: C5 +3 %, the others equal)
#endif
  constexpr int STILE = IU_STILE;
  const int lane = threadIdx.x & 31;
  int64_t t = SCHED ? 0 : (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t tile = 0;
  int sub = STILE / 32;
  int drawn = 0;
  for (;; t += (int64_t)gridDim.x * blockDim.x) {
    if (SCHED) {
      if (sub == STILE / 32) {
        // a bounded number of tiles per warp turns the persistent grid into many short-lived CTAs: the slots they
        // free let the kernels of another stream (the sort of the next sub-batch, api.cu: run_query) run beside this one
        if (P.tiles_per_warp && drawn == P.tiles_per_warp) break;
        ++drawn;
        unsigned long long b = 0;
        if (lane == 0) b = atomicAdd(P.ticket, (unsigned long long)STILE);
        tile = (int64_t)__shfl_sync(0xffffffffu, b, 0);
        sub = 0;
      }
      if (tile >= P.n) break;
      t = tile + 32 * sub + lane;
      ++sub;
      if (t >= P.n) continue;
    } else if (t >= P.n) break;
    const uint32_t pe = P.perm ? LD_STREAM(P.perm + t) : 0u;
    const int64_t i = P.perm ? (int64_t)(pe & ~PERM_HINT_BIT) : t;
    if (IU_OOB(i < 0 || i >= P.n)) continue;
    // a query's 24 bytes as two scattered loads instead of three: 16 + 8 or 8 + 16 bytes, whichever is aligned
    double qx, qy, qz;
    {
      const double* rp = P.r + 3 * i;
      if ((reinterpret_cast<uintptr_t>(rp) & 15) == 0) {
        const double2 a = LD_STREAM(reinterpret_cast<const double2*>(rp));
        qx = a.x; qy = a.y; qz = LD_STREAM(rp + 2);
      } else {
        qx = LD_STREAM(rp);
        const double2 b = LD_STREAM(reinterpret_cast<const double2*>(rp + 1));
        qy = b.x; qz = b.y;
      }
    }
    if (MODE == QM_NEAREST) {
      const int slot = nearest_slot(P, s_lb, qx, qy, qz);
      ST_STREAM(P.out_i32 + i, slot >= 0 ? __ldg(P.orig_of_dev + slot) : 0);
      continue;
    }
    // through a permutation the scattered hint is read only where the key kernel saw one (PERM_HINT_BIT)
    const int h = !P.perm || (pe & PERM_HINT_BIT) ? LD_STREAM(P.i_cell + i) : 0;
    if (h > P.n_cells) { atomicAdd(P.counters + CNT_HINT_ERROR, 1ULL); continue; }     // error stop, :490
    int start;
    if (MODE == QM_INTERP_ICELL) {
      if (h < 1) { atomicAdd(P.counters + CNT_NOT_FOUND, 1ULL); continue; }
      start = __ldg(P.dev_of_orig + h);
    } else if (h < 1) {                                                                // :422-423
      const int slot = nearest_slot(P, s_lb, qx, qy, qz);
      if (slot < 0) { ST_STREAM(P.i_cell + i, 0); atomicAdd(P.counters + CNT_NOT_FOUND, 1ULL); continue; }
      start = slot + 1;
    } else {
      start = __ldg(P.dev_of_orig + h);                                                // :424-426
    }
    finish_query<CT, MODE, NV>(P, i, qx, qy, qz, start);
  }
}

// One iteration of the loop at :698-723 for the lane-refill kernel below: the statements of walk_ray's loop body.
// Returns true while the walk continues (ic1 = the next cell); false when it has ended: ic1 = the containing cell,
// or < 1 where the ray left the domain, or IU_WALK_LIMIT (counted) after max_steps iterations.
template <int PPC>
struct WalkLane { V3 u, r_p; double distance_left; int ic1, n_steps; };

template <int PPC>
__device__ __forceinline__ bool walk_step(const QueryKernelParams& P, WalkLane<PPC>& w) {
  using L = RecLayout<PPC>;
  if (IU_OOB(w.ic1 < 1 || w.ic1 > g_lim[0])) { w.ic1 = 0; return false; }
  const double* R = P.rec + (size_t)(w.ic1 - 1) * L::STRIDE;
  double face_distance = HUGE_DISTANCE;
  int i_face = -1;
  if (PPC == 4) {
#pragma unroll
    for (int h = 0; h < 4; h += 2) {
      V3 fn[2], rf[2];
      ld3n<2>(R + L::NRM + 3 * h, fn);
      ld3n<2>(R + L::PTS + 3 * h, rf);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const double pdn = dot3(w.u, fn[j]);
        const double dist = dot3(rf[j] - w.r_p, fn[j]) / pdn;
        if (pdn > 0 && dist < face_distance) { face_distance = dist; i_face = h + j; }
      }
    }
  } else {
    V3 fn[3], rf[3];
    ld3n<3>(R + L::NRM, fn);
    ld3n<3>(R + L::PTS, rf);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const double pdn = dot3(w.u, fn[k]);
      const double dist = dot3(rf[k] - w.r_p, fn[k]) / pdn;
      if (pdn > 0 && dist < face_distance) { face_distance = dist; i_face = k; }
    }
  }
  w.r_p = {w.r_p.x + face_distance * w.u.x, w.r_p.y + face_distance * w.u.y, w.r_p.z + face_distance * w.u.z};   // :762
  w.distance_left = w.distance_left - face_distance;
  if (!(w.distance_left > 0)) return false;                                            // found, :720-721
  w.ic1 = __ldg(reinterpret_cast<const int32_t*>(R) + L::NBR_I + i_face);
  if (w.ic1 < 1) return false;                                                         // left the domain, :706-709
  if (++w.n_steps > P.max_steps) { w.ic1 = IU_WALK_LIMIT; atomicAdd(P.counters + CNT_WALK_LIMIT, 1ULL); return false; }
  return true;
}

// k_query with LANE REFILL (sorted batches, the modes that walk): a warp draws tiles of STILE consecutive sorted queries
// like k_query<.., SCHED = 1>, but a lane takes the next query of the tile as soon as its own walk has ended.  Every pass
// of the loop = [idle lanes: read the query, nearest centre / hint, set up the ray] + [all lanes: ONE walk step] +
// [lanes whose walk ended: outputs].  A warp of k_query runs ~3.3 walk iterations per 32 queries because some lane always
// needs one more (mean 1.3 per query, 25 % of the lanes in the second iteration, 4 % in the third); here every walk
// step runs with (nearly) all lanes.  Same statements per query in the same order: results are bit-identical.
template <int CT, int MODE, int NV>
__global__ void __launch_bounds__(QUERY_THREADS, IU_QUERY_MIN_BLOCKS) k_query_refill(const QueryKernelParams P) {
  extern __shared__ float s_lb[];
  constexpr int PPC = CT == IU_TRIANGLE_CT ? 3 : 4;
  constexpr int STILE = IU_STILE;
  const int lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u;
  int64_t tile = 0;
  int tile_pos = STILE;         // queries of the current tile already handed out (STILE: draw a new tile)
  bool more = true, busy = false;
  int64_t i = 0;
  double qx = 0, qy = 0, qz = 0;
  WalkLane<PPC> w;
  w.u = w.r_p = V3{0, 0, 0}; w.distance_left = 0; w.ic1 = 0; w.n_steps = 0;
  for (;;) {
    const unsigned idle = __ballot_sync(0xffffffffu, !busy);
    if (idle && more) {
      if (tile_pos >= STILE) {
        unsigned long long b = 0;
        if (lane == 0) b = atomicAdd(P.ticket, (unsigned long long)STILE);
        tile = (int64_t)__shfl_sync(0xffffffffu, b, 0);
        tile_pos = 0;
        if (tile >= P.n) more = false;
      }
      if (more) {
        const int64_t left = P.n - tile;
        const int avail = (int)(left < STILE ? left : STILE) - tile_pos;
        const int rank = __popc(idle & lt);
        const bool take = !busy && rank < avail;
        const int n_idle = __popc(idle);
        const int64_t t = tile + tile_pos + rank;
        tile_pos += n_idle < avail ? n_idle : avail;
        if (avail <= n_idle) tile_pos = STILE;
        if (take) {
          const uint32_t pe = P.perm ? LD_STREAM(P.perm + t) : 0u;
          i = P.perm ? (int64_t)(pe & ~PERM_HINT_BIT) : t;
          bool ok = !IU_OOB(i < 0 || i >= P.n);
          if (ok) {
            const double* rp = P.r + 3 * i;
            if ((reinterpret_cast<uintptr_t>(rp) & 15) == 0) {
              const double2 a = LD_STREAM(reinterpret_cast<const double2*>(rp));
              qx = a.x; qy = a.y; qz = LD_STREAM(rp + 2);
            } else {
              qx = LD_STREAM(rp);
              const double2 b = LD_STREAM(reinterpret_cast<const double2*>(rp + 1));
              qy = b.x; qz = b.y;
            }
            const int h = !P.perm || (pe & PERM_HINT_BIT) ? LD_STREAM(P.i_cell + i) : 0;
            int start = 0;
            if (h > P.n_cells) { atomicAdd(P.counters + CNT_HINT_ERROR, 1ULL); ok = false; }     // error stop, :490
            else if (h < 1) {                                                                    // :422-423
              const int slot = nearest_slot_call(P, s_lb, qx, qy, qz);
              if (slot < 0) { ST_STREAM(P.i_cell + i, 0); atomicAdd(P.counters + CNT_NOT_FOUND, 1ULL); ok = false; }
              start = slot + 1;
            } else start = __ldg(P.dev_of_orig + h);                                             // :424-426
            if (ok && IU_OOB(start < 1 || start > g_lim[0])) ok = false;
            if (ok) {
              const D4 c0 = ld256(P.cen32 + 4 * (size_t)(start - 1));
              const V3 dvec = V3{qx, qy, qz} - V3{c0.x, c0.y, c0.z};                            // r0 = the cell centre, :429
              w.distance_left = norm2_3(dvec);
              w.ic1 = start;
              if (w.distance_left < TINY_DISTANCE) emit_result<CT, MODE, NV>(P, i, qx, qy, qz, start);   // :689-692
              else {
                w.u = V3{dvec.x / w.distance_left, dvec.y / w.distance_left, dvec.z / w.distance_left};  // :694
                w.r_p = V3{c0.x, c0.y, c0.z};
                w.n_steps = 1;
                busy = true;
              }
            }
          }
        }
      }
    }
    if (!__any_sync(0xffffffffu, busy)) {
      if (!more) break;
      continue;
    }
    if (busy && !walk_step<PPC>(P, w)) {
      emit_result<CT, MODE, NV>(P, i, qx, qy, qz, w.ic1);
      busy = false;
    }
  }
}

template <int PPC>
__global__ void __launch_bounds__(QUERY_THREADS) k_walk(const double* __restrict__ rec, const int32_t* __restrict__ dev_of_orig,
                                                        const int32_t* __restrict__ orig_of_dev, const int32_t* __restrict__ mask,
                                                        int n_cells, unsigned long long* counters, WalkArgs a) {
  for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < a.n; t += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = a.perm ? (int64_t)LD_STREAM(a.perm + t) : t;
    if (IU_OOB(i < 0 || i >= a.n)) continue;
    // through a permutation the ray's inputs are scattered reads: streaming loads with 64-byte DRAM fetches, like k_query's
    const int h = a.perm ? LD_STREAM(a.ic0 + i) : a.ic0[i];
    if (h < 1 || h > n_cells) { atomicAdd(counters + CNT_HINT_ERROR, 1ULL); continue; }
    V3 r0, r1;
    if (a.perm) {
      r0 = V3{LD_STREAM(a.r0 + 3 * i), LD_STREAM(a.r0 + 3 * i + 1), LD_STREAM(a.r0 + 3 * i + 2)};
      r1 = V3{LD_STREAM(a.r1 + 3 * i), LD_STREAM(a.r1 + 3 * i + 1), LD_STREAM(a.r1 + 3 * i + 2)};
    } else { r0 = ld3(a.r0 + 3 * i); r1 = ld3(a.r1 + 3 * i); }
    int ic1, n_steps = 0, status; V3 r_p; bool early;
    walk_ray<PPC>(rec, r0, r1, __ldg(dev_of_orig + h), a.max_steps, mask, ic1, r_p, n_steps, status, early);
    a.ic1[i] = ic1 >= 1 ? __ldg(orig_of_dev + (ic1 - 1)) : ic1;
    a.status[i] = status;
    if (!early) {   // r_p / n_steps undefined on the early return, :689-692
      a.r_p[3 * i] = r_p.x; a.r_p[3 * i + 1] = r_p.y; a.r_p[3 * i + 2] = r_p.z;
      a.n_steps[i] = n_steps;
    } else {
      atomicAdd(counters + CNT_EARLY, 1ULL);
      if (a.mark_early) a.n_steps[i] = INT32_MIN;
    }
  }
}

// ------------------------------------------------------------------------------------------------
// iu_integrate_along_field (:989-1217): Bogacki-Shampine RK3(2) along the interpolated field, one
// field line per thread, the reference's statements in the reference's order.  The user callback
// sub_int (:61-74) is replaced by built-in integrands (0 none, 1 dy = 1, 2 dy = norm2(field)).
// The seed's cell and field come from a cold QM_INTERP pass (iu_interpolate_at at :1054).
// ------------------------------------------------------------------------------------------------
__global__ void k_trace_seeds(int64_t n, int ndim, int nt, int max_steps, const double* __restrict__ y, double* __restrict__ r) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double* Y = y + (size_t)i * nt * max_steps;
  r[3 * i] = Y[0];
  r[3 * i + 1] = ndim > 1 ? Y[1] : 0.0;
  r[3 * i + 2] = ndim > 2 ? Y[2] : 0.0;     // "Set unused coordinates to zero", :1045-1047
}

#ifndef IU_TRACE_MIN_BLOCKS
#define IU_TRACE_MIN_BLOCKS 4   // resident CTAs per SM k_trace is compiled for (128 registers; 1 / 4 / 6 / 8 measured: 36.7 / 31.6 / 35.6 / 35.2 ms per 500k lines)
#endif
template <int CT>
__global__ void __launch_bounds__(128, IU_TRACE_MIN_BLOCKS) k_trace(const double* __restrict__ rec, const int32_t* __restrict__ dev_of_orig,
                                               const double* __restrict__ pd, int pd_stride, const int32_t* __restrict__ icd,
                                               int n_cells, unsigned long long* counters, const TraceArgs a) {
  constexpr int PPC = CT == IU_TRIANGLE_CT ? 3 : 4;
  using L = RecLayout<PPC>;
  constexpr int MAXT = 3 + TRACE_MAX_NVAR;
  const double safety_fac = 0.8, inv_24 = 1 / 24.0, inv_9 = 1 / 9.0, min_radius = 1e-12, eps = 1e-8;
  const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= a.n) return;
  // neighbouring threads trace neighbouring lines: their walks read the same records
  const int64_t i = a.perm ? (int64_t)a.perm[t] : t;
  if (IU_OOB(i < 0 || i >= a.n)) return;
  const int ndim = a.ndim, nvar = a.nvar, nt = ndim + nvar;
  double* Y = a.y + (size_t)i * nt * a.max_steps;
  double* YF = a.y_field ? a.y_field + (size_t)i * ndim * a.max_steps : nullptr;
  const int32_t* mask = a.mask_col >= 0 ? icd + (size_t)a.mask_col * n_cells : nullptr;

  int n_steps = 1;
  const int ic_orig = a.i_cell0[i];
  const bool valid = ic_orig > 0 && (!mask || __ldg(mask + (ic_orig - 1)) == a.mask_value);      // cell_is_valid, :1203-1215
  if (!valid) {                                                                                   // :1057-1070
    if (a.boundary_material) a.boundary_material[i] = ic_orig <= 0 ? -1 : (mask ? __ldg(mask + (ic_orig - 1)) : 0);
    a.n_steps[i] = n_steps;
    return;
  }
  double yc[MAXT], ys[MAXT], y2[MAXT], k[4][MAXT], field[3], fprev[3];
#pragma unroll
  for (int c = 0; c < MAXT; ++c) { yc[c] = c < nt ? Y[c] : 0.0; ys[c] = 0; y2[c] = 0; k[0][c] = k[1][c] = k[2][c] = k[3][c] = 0; }
#pragma unroll
  for (int c = 0; c < 3; ++c) { fprev[c] = c < ndim ? a.field0[(size_t)i * ndim + c] : 0.0; field[c] = fprev[c]; }
  if (YF) for (int c = 0; c < ndim; ++c) YF[c] = fprev[c];
  int ic_prev = __ldg(dev_of_orig + ic_orig), ic = ic_prev, status = 0;
  long long last_rejected = -100;
  double dx = a.max_dx;
  V3 r0{0, 0, 0}, r_p{0, 0, 0};

  auto unitvec = [&](double* out) {                                                              // get_unitvec, :1192-1201
    const double nrm = norm2_3(V3{field[0], field[1], field[2]});
#pragma unroll
    for (int c = 0; c < 3; ++c) if (c < ndim) { out[c] = field[c] / nrm; if (a.reverse) out[c] = -out[c]; }
    double dv = 1.0;
    if (a.integrand == 2) dv = nrm;
    for (int c = 0; c < nvar; ++c) out[ndim + c] = dv;                                           // sub_int
  };
  // position ys -> r, walk from (r0, ic_prev), interpolate the field there; false => "cycle"
  auto substep = [&](double* kout) -> bool {
    V3 r{ys[0], ndim > 1 ? ys[1] : 0.0, ndim > 2 ? ys[2] : 0.0};
    if (a.axisymmetric) r.x = fmax(r.x, min_radius);
    int n_its = 0; bool early;
    walk_ray<PPC>(rec, r0, r, ic_prev, a.max_walk_steps, mask, ic, r_p, n_its, status, early);
    if (!early && n_its > a.max_walk_steps) { status = -1; ic = 0; atomicAdd(counters + CNT_WALK_LIMIT, 1ULL); }
    if (status != 0) return false;
    const double* R = rec + (size_t)(ic - 1) * L::STRIDE;                                         // iu_interpolate_at_icell
    const int32_t* RI = reinterpret_cast<const int32_t*>(R);
    V3 p[PPC];
    ld3n<PPC>(R + L::PTS, p);
    Weights<CT> W;
    W.setup(p, __ldg(R + L::VOL), r);
    for (int v = 0; v < ndim; ++v) {
      double d[4];
#pragma unroll
      for (int q = 0; q < PPC; ++q) d[q] = __ldg(pd + (size_t)__ldg(RI + L::CEL_I + q) * pd_stride + a.ivars[v]);
      field[v] = W.apply(d);
    }
    unitvec(kout);
    return true;
  };
  auto boundary = [&]() {                                                                         // :1087-1093
    if (!a.boundary_material) return;
    int bm = -1;
    if (ic > 0 && mask) bm = __ldg(mask + (__ldg(reinterpret_cast<const int32_t*>(rec + (size_t)(ic - 1) * L::STRIDE) + L::ORG_I) - 1));
    a.boundary_material[i] = bm;
  };

  const long long iter_cap = 64LL * a.max_steps + 100000;   // the reference loops to huge(1)-1; every pass shrinks dx or advances
  for (long long iteration = 1; iteration <= iter_cap; ++iteration) {
    if (status != 0) {                                                                            // :1081-1097
      last_rejected = iteration - 1;
      dx = (1 - eps) * norm2_3(r_p - r0);
      if (dx < a.min_dx) { boundary(); break; }
    }
    ic = ic_prev;
    r0 = V3{yc[0], ndim > 1 ? yc[1] : 0.0, ndim > 2 ? yc[2] : 0.0};
#pragma unroll
    for (int c = 0; c < 3; ++c) field[c] = fprev[c];
    unitvec(k[0]);
    { const double f = 0.5 * dx;
#pragma unroll
      for (int c = 0; c < MAXT; ++c) if (c < nt) ys[c] = yc[c] + f * k[0][c]; }                    // :1118
    if (!substep(k[1])) continue;
    { const double f = 0.75 * dx;
#pragma unroll
      for (int c = 0; c < MAXT; ++c) if (c < nt) ys[c] = yc[c] + f * k[1][c]; }                    // :1131
    if (!substep(k[2])) continue;
    { const double f = dx * inv_9;
#pragma unroll
      for (int c = 0; c < MAXT; ++c) if (c < nt) ys[c] = yc[c] + f * ((2 * k[0][c] + 3 * k[1][c]) + 4 * k[2][c]); }   // :1144-1145
    if (!substep(k[3])) continue;
    double sum = 0;
    { const double f = dx * inv_24;
#pragma unroll
      for (int c = 0; c < MAXT; ++c) if (c < nt) {
        y2[c] = yc[c] + f * (((7 * k[0][c] + 6 * k[1][c]) + 8 * k[2][c]) + 3 * k[3][c]);         // :1159-1160
        const double scale = a.atol + fmax(fabs(ys[c]), fabs(y2[c])) * a.rtol;
        const double e = (ys[c] - y2[c]) / scale;
        sum += e * e;
      } }
    const double err = sqrt(sum / 3);
    if (err <= 1 || dx < 2 * a.min_dx) {                                                          // :1165-1176
      n_steps = n_steps + 1;
      if (n_steps > a.max_steps) break;
#pragma unroll
      for (int c = 0; c < MAXT; ++c) if (c < nt) yc[c] = ys[c];
      if (a.axisymmetric) yc[0] = fmax(yc[0], min_radius);
      for (int c = 0; c < nt; ++c) Y[(size_t)(n_steps - 1) * nt + c] = yc[c];
#pragma unroll
      for (int c = 0; c < 3; ++c) fprev[c] = field[c];
      if (YF) for (int c = 0; c < ndim; ++c) YF[(size_t)(n_steps - 1) * ndim + c] = field[c];
      ic_prev = ic;
    } else {
      last_rejected = iteration;
    }
    const double max_growth = last_rejected > iteration - 2 ? 1.0 : 2.0;                          // :1178-1184
    const double dx_factor = fmin(max_growth, safety_fac * pow(1 / err, 1 / 3.0));
    dx = fmax(a.min_dx, fmin(a.max_dx, dx * dx_factor));
    if (iteration == iter_cap) atomicAdd(counters + CNT_WALK_LIMIT, 1ULL);
  }
  a.n_steps[i] = n_steps;
}

int sm_count() {
  static int n_sm = 0;
  if (n_sm == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    if (n_sm <= 0) n_sm = 148;
  }
  return n_sm;
}

// persistent grid: SMs x resident CTAs for this kernel and dynamic shared memory size
int query_grid_blocks(const void* kernel, size_t dyn_smem) {
  int per_sm = 0;
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, QUERY_THREADS, dyn_smem);
  if (per_sm < 1) per_sm = 1;
  return sm_count() * per_sm;
}

template <int CT, int MODE, int NV, int SCHED>
void launch_query_static(const QueryKernelParams& P, int depth, cudaStream_t s) {
  const size_t dyn = MODE == QM_INTERP_ICELL ? 0 : (size_t)(depth > 0 ? depth : 1) * QUERY_THREADS * sizeof(float);
  static int blocks = 0;
  static size_t blocks_dyn = ~(size_t)0;
  if (blocks == 0 || blocks_dyn != dyn) {
    if (dyn + 20 * 1024 > 48 * 1024)
      cudaFuncSetAttribute((const void*)k_query<CT, MODE, NV, SCHED>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
    blocks = query_grid_blocks((const void*)k_query<CT, MODE, NV, SCHED>, dyn);
    blocks_dyn = dyn;
  }
  int64_t need = (P.n + QUERY_THREADS - 1) / QUERY_THREADS;
  int g = (int)(need < blocks ? need : blocks);
  if (SCHED && P.tiles_per_warp > 0) {   // every tile must find a warp: tiles / (warps per CTA x tiles per warp) CTAs
    const int64_t tiles = (P.n + IU_STILE - 1) / IU_STILE;
    const int64_t per_cta = (int64_t)(QUERY_THREADS / 32) * P.tiles_per_warp;
    g = (int)std::max<int64_t>(1, (tiles + per_cta - 1) / per_cta);
  }
  if (SCHED) cudaMemsetAsync(P.ticket, 0, sizeof(unsigned long long), s);
  k_query<CT, MODE, NV, SCHED><<<g, QUERY_THREADS, dyn, s>>>(P);
}

template <int CT, int MODE, int NV>
void launch_query_refill(const QueryKernelParams& P, int depth, cudaStream_t s) {
  const size_t dyn = (size_t)(depth > 0 ? depth : 1) * QUERY_THREADS * sizeof(float);
  static int blocks = 0;
  static size_t blocks_dyn = ~(size_t)0;
  if (blocks == 0 || blocks_dyn != dyn) {
    if (dyn + 20 * 1024 > 48 * 1024)
      cudaFuncSetAttribute((const void*)k_query_refill<CT, MODE, NV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
    blocks = query_grid_blocks((const void*)k_query_refill<CT, MODE, NV>, dyn);
    blocks_dyn = dyn;
  }
  const int64_t need = (P.n + QUERY_THREADS - 1) / QUERY_THREADS;
  const int g = (int)(need < blocks ? need : blocks);
  cudaMemsetAsync(P.ticket, 0, sizeof(unsigned long long), s);
  k_query_refill<CT, MODE, NV><<<g, QUERY_THREADS, dyn, s>>>(P);
}

// sched: 0 static grid-stride, 1 ticket tiles (sorted batches), 2 ticket tiles with lane refill (the modes that walk)
template <int CT, int MODE, int NV>
void launch_query_inst(const QueryKernelParams& P, int depth, cudaStream_t s) {
  constexpr bool walks = MODE == QM_CELL || MODE == QM_INTERP || MODE == QM_CELL_F64 || MODE == QM_CELL_I32 || MODE == QM_INTERP_CELL;
  if constexpr (walks) {
    if (P.ticket && P.sched == 2) { launch_query_refill<CT, MODE, NV>(P, depth, s); return; }
  }
  if (P.ticket && P.sched >= 1) launch_query_static<CT, MODE, NV, 1>(P, depth, s);
  else launch_query_static<CT, MODE, NV, 0>(P, depth, s);
}

template <int CT>
void launch_query_ct(int mode, const QueryKernelParams& P, int depth, cudaStream_t s) {
  switch (mode) {
    case QM_NEAREST: launch_query_inst<CT, QM_NEAREST, 0>(P, depth, s); break;
    case QM_CELL: launch_query_inst<CT, QM_CELL, 0>(P, depth, s); break;
    case QM_CELL_F64: launch_query_inst<CT, QM_CELL_F64, 0>(P, depth, s); break;
    case QM_CELL_I32: launch_query_inst<CT, QM_CELL_I32, 0>(P, depth, s); break;
    case QM_INTERP:
      switch (P.n_vars) {
        case 1: launch_query_inst<CT, QM_INTERP, 1>(P, depth, s); break;
        case 2: launch_query_inst<CT, QM_INTERP, 2>(P, depth, s); break;
        case 3: launch_query_inst<CT, QM_INTERP, 3>(P, depth, s); break;
        case 4: launch_query_inst<CT, QM_INTERP, 4>(P, depth, s); break;
        case 8: launch_query_inst<CT, QM_INTERP, 8>(P, depth, s); break;
        default: launch_query_inst<CT, QM_INTERP, 0>(P, depth, s); break;
      }
      break;
    case QM_INTERP_CELL:
      switch (P.n_vars) {
        case 4: launch_query_inst<CT, QM_INTERP_CELL, 4>(P, depth, s); break;
        case 8: launch_query_inst<CT, QM_INTERP_CELL, 8>(P, depth, s); break;
        default: launch_query_inst<CT, QM_INTERP_CELL, 0>(P, depth, s); break;
      }
      break;
    case QM_INTERP_ICELL:
      switch (P.n_vars) {
        case 1: launch_query_inst<CT, QM_INTERP_ICELL, 1>(P, depth, s); break;
        case 2: launch_query_inst<CT, QM_INTERP_ICELL, 2>(P, depth, s); break;
        case 3: launch_query_inst<CT, QM_INTERP_ICELL, 3>(P, depth, s); break;
        case 4: launch_query_inst<CT, QM_INTERP_ICELL, 4>(P, depth, s); break;
        default: launch_query_inst<CT, QM_INTERP_ICELL, 0>(P, depth, s); break;
      }
      break;
  }
}

inline int blocks_for(int64_t n, int threads) { return (int)((n + threads - 1) / threads); }

}  // namespace

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
int launch_cell_centers(const DeviceGrid& g, const double* points, const int32_t* cells, double* centers, cudaStream_t s) {
  if (g.n_cells == 0) return 0;
  const int b = blocks_for(g.n_cells, 256);
  if (g.ppc == 3) k_cell_centers<3><<<b, 256, 0, s>>>(g.n_cells, points, cells, centers);
  else k_cell_centers<4><<<b, 256, 0, s>>>(g.n_cells, points, cells, centers);
  return 1;
}

int launch_build_records(const DeviceGrid& g, const double* points, const int32_t* cells, const int32_t* neighbors,
                         const double* centers, cudaStream_t s) {
  const int n_slots = g.n_leaves * LEAF;
  if (n_slots == 0) return 0;
  const int b = blocks_for(n_slots, 128);
  k_fill_dev_of_orig<<<blocks_for(n_slots, 256), 256, 0, s>>>(n_slots, g.orig_of_dev, g.dev_of_orig);
  if (g.cell_type == IU_TRIANGLE_CT)
    k_build_records<3, IU_TRIANGLE_CT><<<b, 128, 0, s>>>(n_slots, g.orig_of_dev, g.dev_of_orig, points, cells, neighbors, centers, g.rec, g.centers4, g.cen32);
  else if (g.cell_type == IU_QUAD_CT)
    k_build_records<4, IU_QUAD_CT><<<b, 128, 0, s>>>(n_slots, g.orig_of_dev, g.dev_of_orig, points, cells, neighbors, centers, g.rec, g.centers4, g.cen32);
  else
    k_build_records<4, IU_TETRA_CT><<<b, 128, 0, s>>>(n_slots, g.orig_of_dev, g.dev_of_orig, points, cells, neighbors, centers, g.rec, g.centers4, g.cen32);
  return 2;
}

int launch_export_geometry(const DeviceGrid& g, double* cell_points, double* normals, double* volume, double* centers,
                           cudaStream_t s) {
  const int n_slots = g.n_leaves * LEAF;
  if (n_slots == 0) return 0;
  const int b = blocks_for(n_slots, 128);
  if (g.ppc == 3) k_export_geometry<3><<<b, 128, 0, s>>>(n_slots, g.orig_of_dev, g.rec, g.centers4, cell_points, normals, volume, centers);
  else k_export_geometry<4><<<b, 128, 0, s>>>(n_slots, g.orig_of_dev, g.rec, g.centers4, cell_points, normals, volume, centers);
  return 1;
}

int launch_export_mesh(const DeviceGrid& g, double* points, int32_t* cells, cudaStream_t s) {
  const int n_slots = g.n_leaves * LEAF;
  if (n_slots == 0) return 0;
  const int b = blocks_for(n_slots, 128);
  if (g.ppc == 3) k_export_mesh<3><<<b, 128, 0, s>>>(n_slots, g.orig_of_dev, g.rec, points, cells);
  else k_export_mesh<4><<<b, 128, 0, s>>>(n_slots, g.orig_of_dev, g.rec, points, cells);
  return 1;
}

int launch_pack_point_data(const DeviceGrid& g, const double* cols, int col0, int n_cols, cudaStream_t s) {
  const int64_t n = (int64_t)g.n_points * n_cols;
  if (n == 0) return 0;
  k_pack_point_data<<<blocks_for(n, 256), 256, 0, s>>>(g.n_points, n_cols, col0, g.pd_stride, cols, g.pd);
  return 1;
}

int launch_unpack_point_data(const DeviceGrid& g, double* cols, int col0, int n_cols, cudaStream_t s) {
  const int64_t n = (int64_t)g.n_points * n_cols;
  if (n == 0) return 0;
  k_unpack_point_data<<<blocks_for(n, 256), 256, 0, s>>>(g.n_points, n_cols, col0, g.pd_stride, g.pd, cols);
  return 1;
}

int launch_repack_point_data(const double* old_pd, int old_stride, double* new_pd, int new_stride, int n_cols, int n_points,
                             cudaStream_t s) {
  const int64_t n = (int64_t)n_points * new_stride;
  if (n == 0) return 0;
  k_repack_point_data<<<blocks_for(n, 256), 256, 0, s>>>(n_points, n_cols, old_stride, new_stride, old_pd, new_pd);
  return 1;
}

int launch_generate_queries(int64_t n, int64_t first, uint64_t seed, const double lo[3], const double hi[3], double* r,
                            cudaStream_t s) {
  if (n == 0) return 0;
  k_generate_queries<<<blocks_for(3 * n, 256), 256, 0, s>>>(n, first, seed, lo[0], lo[1], lo[2], hi[0], hi[1], hi[2], r);
  return 1;
}

int launch_merge_rows(int64_t n, int row_bytes, const int32_t* i_cell, const void* src, void* dst, cudaStream_t s) {
  if (n == 0) return 0;
  const int words = row_bytes / 4;
  k_merge_rows<<<blocks_for(n * words, 256), 256, 0, s>>>(n, words, i_cell, (const uint32_t*)src, (uint32_t*)dst);
  return 1;
}

// bits per dimension of the Morton key: about one key cell per mesh cell
static void morton_bits(const DeviceGrid& g, int& bits, int& planar) {
  planar = (g.rmax[2] - g.rmin[2]) == 0.0 ? 1 : 0;
  double lg = 0;
  for (int64_t c = 1; c < g.n_cells; c <<= 1) lg += 1;
  if (planar) { bits = (int)((lg + 1) / 2); bits = bits < 1 ? 1 : (bits > 15 ? 15 : bits); }
  else { bits = (int)((lg + 2) / 3); bits = bits < 1 ? 1 : (bits > 10 ? 10 : bits); }
}

int morton_end_bit(const DeviceGrid& g) {
  int bits, planar;
  morton_bits(g, bits, planar);
  return planar ? 2 * bits : 3 * bits;
}

size_t sort_temp_bytes(int64_t n, int end_bit) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr,
                                  (uint32_t*)nullptr, n, 0, end_bit);
  return bytes;
}

int launch_morton_sort(const DeviceGrid& g, int64_t n, const double* r, const int32_t* hint, uint32_t* keys, uint32_t* keys_alt,
                       uint32_t* idx, uint32_t* idx_alt, void* temp, size_t temp_bytes, int* end_bit_out, cudaStream_t s) {
  int bits, planar;
  morton_bits(g, bits, planar);
  const double nb = (double)(1u << bits);
  double sc[3];
  for (int k = 0; k < 3; ++k) { const double ext = g.rmax[k] - g.rmin[k]; sc[k] = ext > 0 ? nb / ext : 0.0; }
  k_morton_keys<<<(unsigned)std::min<int64_t>(blocks_for(n, 256), 148 * 32), 256, 0, s>>>(n, r, hint, g.rmin[0], g.rmin[1], g.rmin[2], sc[0], sc[1], sc[2], bits, planar, keys, idx);
  const int end_bit = planar ? 2 * bits : 3 * bits;
  // experiment knob: ignore the lowest key bits (queries of one coarse key cell keep their input order)
  static const int drop_bits = [] { const char* e = std::getenv("IU_B200_SORT_DROP_BITS"); return e ? std::atoi(e) : 0; }();
  const int begin_bit = drop_bits > 0 && drop_bits < end_bit ? drop_bits : 0;
  cub::DeviceRadixSort::SortPairs(temp, temp_bytes, (const uint32_t*)keys, keys_alt, (const uint32_t*)idx, idx_alt, n, begin_bit, end_bit, s);
  if (end_bit_out) *end_bit_out = end_bit;
  // cub one-sweep: 1 histogram + 1 scan-ish + ceil(end_bit/8) passes; counted conservatively
  return 1 + 2 + (end_bit + 7) / 8;
}

int launch_build_voxels(const DeviceGrid& g, const VoxelGrid& v, int64_t n_vox, uint32_t* lists, unsigned long long* stats,
                        cudaStream_t s) {
  if (n_vox == 0) return 0;
  const TreeView t{g.fat, g.centers4, g.orig_of_dev, g.box, g.n_internal};
  const int64_t n_blocks = (int64_t)((v.n[0] + 1) / 2) * ((v.n[1] + 1) / 2) * ((v.n[2] + 1) / 2);   // 2 x 2 x 2 voxels per warp
  const int64_t need = (n_blocks + 7) / 8;
  const int b = (int)std::min<int64_t>(need, (int64_t)sm_count() * 8);
  k_build_voxels<<<b, 256, 0, s>>>(t, v, n_vox, lists, stats);
  return 1;
}

// checked builds: the index limits of this grid for the kernels about to be launched on s
static void set_check_limits(const DeviceGrid& g, cudaStream_t s) {
#if IU_CHECKED
  const long long lim[4] = {(long long)g.n_leaves * LEAF, g.n_points, (long long)g.n_vox, g.n_leaves};
  cudaMemcpyToSymbolAsync(g_lim, lim, sizeof(lim), 0, cudaMemcpyHostToDevice, s);
#else
  (void)g; (void)s;
#endif
}

long long bounds_errors() {
#if IU_CHECKED
  unsigned long long v = 0;
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(&v, g_bounds_errors, sizeof(v));
  return (long long)v;
#else
  return -1;
#endif
}

int launch_query(const DeviceGrid& g, const QueryArgs& a, cudaStream_t s) {
  if (a.n == 0) return 0;
  set_check_limits(g, s);
  QueryKernelParams P;
  P.tree = TreeView{g.fat, g.centers4, g.orig_of_dev, g.box, g.n_internal};
  P.vox = g.vox; P.cen32 = g.cen32;
  P.rec = g.rec; P.dev_of_orig = g.dev_of_orig; P.orig_of_dev = g.orig_of_dev;
  P.pd = g.pd; P.cd = g.cd; P.icd = g.icd; P.counters = g.counters;
  P.r = a.r; P.perm = a.perm; P.i_cell = a.i_cell; P.out_f64 = a.out_f64; P.out_i32 = a.out_i32;
  P.out2_f64 = a.out2_f64; P.cell_var = a.cell_var;
  P.n = a.n; P.n_cells = g.n_cells; P.pd_stride = g.pd_stride; P.n_vars = a.n_vars; P.max_steps = a.max_steps;
  for (int v = 0; v < MAX_VARS; ++v) P.ivars[v] = a.ivars[v];
  P.pairs = 0;
  if ((a.mode == QM_INTERP || a.mode == QM_INTERP_ICELL || a.mode == QM_INTERP_CELL) && a.n_vars % 2 == 0 && g.pd_stride % 2 == 0) {
    P.pairs = 1;
    for (int v = 0; v < a.n_vars; v += 2)
      if (a.ivars[v] % 2 != 0 || a.ivars[v + 1] != a.ivars[v] + 1) P.pairs = 0;
    if (P.pairs && a.n_vars % 4 == 0 && g.pd_stride % 4 == 0) {
      P.pairs = 2;
      for (int v = 0; v < a.n_vars; v += 4)
        if (a.ivars[v] % 4 != 0 || a.ivars[v + 2] != a.ivars[v] + 2) P.pairs = 1;
    }
  }
  P.ticket = a.dynamic ? (a.ticket ? a.ticket : g.counters + CNT_TICKET) : nullptr;
  P.sched = a.dynamic;
  P.tiles_per_warp = a.tiles_per_warp;
  if (g.cell_type == IU_TRIANGLE_CT) launch_query_ct<IU_TRIANGLE_CT>(a.mode, P, g.depth, s);
  else if (g.cell_type == IU_QUAD_CT) launch_query_ct<IU_QUAD_CT>(a.mode, P, g.depth, s);
  else launch_query_ct<IU_TETRA_CT>(a.mode, P, g.depth, s);
  return 1;
}

int launch_trace_seeds(int64_t n, int ndim, int nt, int max_steps, const double* y, double* r, cudaStream_t s) {
  if (n == 0) return 0;
  k_trace_seeds<<<blocks_for(n, 256), 256, 0, s>>>(n, ndim, nt, max_steps, y, r);
  return 1;
}

int launch_trace(const DeviceGrid& g, const TraceArgs& a, cudaStream_t s) {
  if (a.n == 0) return 0;
  set_check_limits(g, s);
  const int b = blocks_for(a.n, 128);
  if (g.cell_type == IU_TRIANGLE_CT)
    k_trace<IU_TRIANGLE_CT><<<b, 128, 0, s>>>(g.rec, g.dev_of_orig, g.pd, g.pd_stride, g.icd, g.n_cells, g.counters, a);
  else if (g.cell_type == IU_QUAD_CT)
    k_trace<IU_QUAD_CT><<<b, 128, 0, s>>>(g.rec, g.dev_of_orig, g.pd, g.pd_stride, g.icd, g.n_cells, g.counters, a);
  else
    k_trace<IU_TETRA_CT><<<b, 128, 0, s>>>(g.rec, g.dev_of_orig, g.pd, g.pd_stride, g.icd, g.n_cells, g.counters, a);
  return 1;
}

// sum of |r1 - r0| over n_samples rays spread evenly over the batch (run_walk's sort decision)
__global__ void k_ray_length_sample(int64_t n, int n_samples, const double* __restrict__ r0, const double* __restrict__ r1,
                                    double* __restrict__ sum) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  double len = 0.0;
  if (j < n_samples) {
    const int64_t i = (int64_t)((double)j * (double)n / (double)n_samples);
    const double dx = r1[3 * i] - r0[3 * i], dy = r1[3 * i + 1] - r0[3 * i + 1], dz = r1[3 * i + 2] - r0[3 * i + 2];
    len = sqrt(dx * dx + dy * dy + dz * dz);
    if (!(len < 1.0e300)) len = 0.0;   // NaN / inf rays do not vote
  }
  for (int o = 16; o > 0; o >>= 1) len += __shfl_down_sync(0xffffffffu, len, o);
  if ((threadIdx.x & 31) == 0) atomicAdd(sum, len);
}

int launch_ray_length_sample(int64_t n, int n_samples, const double* r0, const double* r1, double* sum, cudaStream_t s) {
  cudaMemsetAsync(sum, 0, sizeof(double), s);
  k_ray_length_sample<<<(n_samples + 255) / 256, 256, 0, s>>>(n, n_samples, r0, r1, sum);
  return 1;
}

// ------------------------------------------------------------------------------------------------
// The allocation-free query pass (iu_b200_interpolate_at_batch_dev_noalloc; no counterpart in the reference, added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops).
// Scratch layout inside g_noalloc_scratch, all offsets multiples of 256 bytes:
//   [0, 256)            ticket counter of the tile scheduler (cleared by every call)
//   4 x 4 MiB           curve keys, sorted keys, query indices, permutation for up to NOALLOC_MAX_QUERIES queries
//   rest (8 MiB)        temporary storage of cub::DeviceRadixSort::SortPairs
// Every GPU that loads the library holds the 24 MiB + 256 B until exit.  Every call shares the array: calls that overlap
// on the GPU (different streams, different grids) race on it.
// ------------------------------------------------------------------------------------------------
constexpr size_t NOALLOC_ARR = (size_t)NOALLOC_MAX_QUERIES * 4;
constexpr size_t NOALLOC_TEMP = (size_t)8 << 20;
constexpr size_t NOALLOC_BYTES = 256 + 4 * NOALLOC_ARR + NOALLOC_TEMP;
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_BYTES];

// In host code g_noalloc_scratch is only a stand-in for the symbol: its device address is looked up once per device,
// with that device current.
static unsigned char* noalloc_scratch_ptr() {
  static unsigned char* ptr[64] = {nullptr};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) { set_error("noalloc: cannot identify the current device"); return nullptr; }
  if (!ptr[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) { set_error(std::string("noalloc: cudaGetSymbolAddress: ") + cudaGetErrorString(e)); return nullptr; }
    ptr[dev] = (unsigned char*)p;
  }
  return ptr[dev];
}

int launch_query_noalloc(const DeviceGrid& g, QueryArgs a, cudaStream_t s) {
  if (a.n == 0) return 0;
  if (a.n < 0 || a.n > NOALLOC_MAX_QUERIES) { set_error("noalloc: more than 1048576 queries in one call"); return -1; }
  unsigned char* base = noalloc_scratch_ptr();
  if (!base) return -1;
  unsigned long long* ticket = reinterpret_cast<unsigned long long*>(base);
  uint32_t* keys = reinterpret_cast<uint32_t*>(base + 256);
  uint32_t* keys_alt = reinterpret_cast<uint32_t*>(base + 256 + NOALLOC_ARR);
  uint32_t* idx = reinterpret_cast<uint32_t*>(base + 256 + 2 * NOALLOC_ARR);
  uint32_t* perm = reinterpret_cast<uint32_t*>(base + 256 + 3 * NOALLOC_ARR);
  void* temp = base + 256 + 4 * NOALLOC_ARR;
  const size_t need = sort_temp_bytes(a.n, morton_end_bit(g));   // a size query: allocates nothing
  if (need > NOALLOC_TEMP) { set_error("noalloc: the radix sort wants more temporary storage than the static scratch reserves"); return -1; }
  int launches = launch_morton_sort(g, a.n, a.r, a.mode == QM_NEAREST ? nullptr : a.i_cell, keys, keys_alt, idx, perm, temp, NOALLOC_TEMP,
                                    nullptr, s);
  a.perm = perm;
  a.dynamic = 1;
  a.tiles_per_warp = 0;
  a.ticket = ticket;       // launch_query_static clears it (cudaMemsetAsync on the stream) before the kernel starts
  launches += launch_query(g, a, s);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { set_error(std::string("noalloc: ") + cudaGetErrorString(e)); return -1; }
  return launches;
}

int launch_walk(const DeviceGrid& g, const WalkArgs& a, cudaStream_t s) {
  if (a.n == 0) return 0;
  set_check_limits(g, s);
  const int32_t* mask = a.mask_col >= 0 ? g.icd + (size_t)a.mask_col * g.n_cells : nullptr;
  const int64_t need = (a.n + QUERY_THREADS - 1) / QUERY_THREADS;
  const int b = (int)(need < sm_count() * 8 ? need : sm_count() * 8);
  if (g.ppc == 3) k_walk<3><<<b, QUERY_THREADS, 0, s>>>(g.rec, g.dev_of_orig, g.orig_of_dev, mask, g.n_cells, g.counters, a);
  else k_walk<4><<<b, QUERY_THREADS, 0, s>>>(g.rec, g.dev_of_orig, g.orig_of_dev, mask, g.n_cells, g.counters, a);
  return 1;
}

}  // namespace iu
