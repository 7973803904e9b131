// device.h -- launch interface between api.cu (C ABI, pipelines) and kernels.cu (sm_100a kernels)
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "internal.h"

namespace iu {

// Per-cell record, device order (= kd-tree leaf-slot order), in doubles:
//   ppc = 4 (quad, tetra), 256 B: [0..11] face normals n1..n4, [12..23] vertices p1..p4,
//       [24,25] 4 x int32 vertex ids (0-based), [26] volume, [27] int32 original id (1-based) + pad
//       (one aligned 32-byte sector = everything the interpolation needs besides the vertices),
//       [28,29] 4 x int32 neighbours (1-based device ids, <= 0 passed through)
//   ppc = 3 (triangle), 192 B: [0..8] normals, [9] area, [10..18] vertices, [19] pad,
//       [20,21] int32 {nb1, nb2, nb3, original id}, [22,23] int32 {c1, c2, c3, pad}
template <int PPC> struct RecLayout;
template <> struct RecLayout<4> {
  static constexpr int STRIDE = 32, NRM = 0, PTS = 12, CEL_I = 48, VOL = 26, ORG_I = 54, NBR_I = 56, TAIL = 24;
};
template <> struct RecLayout<3> {
  static constexpr int STRIDE = 24, NRM = 0, PTS = 10, NBR_I = 40, CEL_I = 44, VOL = 9, ORG_I = 43;
};
inline int rec_stride(int ppc) { return ppc == 3 ? RecLayout<3>::STRIDE : RecLayout<4>::STRIDE; }

enum QueryMode : int {
  QM_NEAREST = 0,       // find_nearby_cell_kdtree            -> out_i32 = nearest cell
  QM_CELL = 1,          // iu_get_cell                         -> i_cell
  QM_INTERP = 2,        // iu_interpolate_at                   -> i_cell, out_f64 (n_vars per query)
  QM_CELL_F64 = 3,      // iu_get_cell_scalar_at               -> i_cell, out_f64
  QM_CELL_I32 = 4,      // iu_get_icell_scalar_at              -> i_cell, out_i32
  QM_INTERP_ICELL = 5,  // iu_interpolate_at_icell (cells given, no search)
  QM_INTERP_CELL = 6    // iu_interpolate_at followed by iu_get_cell_scalar_at with the found cell as hint, in one pass
                        //                                     -> i_cell, out_f64 (n_vars per query), out2_f64
};

// counters (device, unsigned long long each)
enum { CNT_HINT_ERROR = 0, CNT_WALK_LIMIT = 1, CNT_NOT_FOUND = 2, CNT_EARLY = 3, CNT_TICKET = 4, CNT_NN_TREE = 5, CNT_N = 6 };

// Voxel table of the nearest-centre search (kernels.cu: k_build_voxels / nn_voxel).  A regular grid over the bounding box
// of the points; every voxel owns VOX_K slot ids = a SUPERSET of all centres that are the nearest centre of some point of
// the voxel (its Voronoi cells, found with the kd-tree at set-up).  A query inside the grid evaluates those candidates
// with the exact fp64 distance of the tree's leaf scan; voxels with more than VOX_K candidates carry VOX_TREE and
// queries outside the grid use the kd-tree.
constexpr int VOX_K = 16;
constexpr uint32_t VOX_END = 0xffffffffu, VOX_TREE = 0xfffffffeu;
struct VoxelGrid {
  const uint32_t* lists = nullptr;   // n_vox x VOX_K slot ids (groups of four; a group starting with VOX_END ends the list)
  double lo[3] = {0, 0, 0};          // origin = rmin of the points
  double inv[3] = {0, 0, 0};         // voxels per unit length (0 in flat dimensions)
  double h[3] = {0, 0, 0};           // voxel size (0 in flat dimensions)
  double flat_val[3] = {0, 0, 0};    // flat dimension: the coordinate every centre has there (queries must match it exactly)
  int n[3] = {1, 1, 1};
  int flat_mask = 0;                 // bit k: dimension k has zero extent
};

struct DeviceGrid {
  int cell_type = 0, ppc = 0, n_cells = 0, n_points = 0;
  // kd-tree
  int n_leaves = 0, n_internal = 0;
  int depth = 0;                   // levels of internal nodes = level of the deepest leaf
  double* fat = nullptr;           // n_internal x {split, left child's split, right child's split, bits(3 x 2-bit dimension)}:
                                   // one 32-byte load decides two levels of a descent
  double* centers4 = nullptr;      // n_leaves x {x[LEAF], y[LEAF], z[LEAF]} (192-byte leaf blocks); padding slots = +inf
  double* cen32 = nullptr;         // n_leaves*LEAF x {x, y, z, original id (low word)}: one 32-byte load per candidate
  VoxelGrid vox;                   // lists == nullptr: no table, every cold query searches the tree
  uint32_t* vox_lists = nullptr;   // owning pointer behind vox.lists
  int64_t n_vox = 0;
  float* box = nullptr;            // (n_internal + n_leaves) x 8 floats: tight box of the centres below each node
  // mesh
  double* rec = nullptr;           // n_leaves*LEAF records
  double* points = nullptr;        // (3, n_points) as given: iu_write_vtk writes ug%points, including points no cell uses
  int32_t* dev_of_orig = nullptr;  // [n_cells + 1], index = 1-based original id -> 1-based device id
  int32_t* orig_of_dev = nullptr;  // [n_leaves*LEAF], 1-based original id, 0 = padding
  // data tables
  int n_pd = 0, pd_stride = 0;
  double* pd = nullptr;            // point-major [n_points][pd_stride]
  int n_cd = 0, cd_cap = 0;
  double* cd = nullptr;            // column-major (n_cells, n_cd), original cell order
  int n_icd = 0, icd_cap = 0;
  int32_t* icd = nullptr;          // column-major (n_cells, n_icd), original cell order
  unsigned long long* counters = nullptr;
  double rmin[3] = {0, 0, 0}, rmax[3] = {0, 0, 0};
};

struct QueryArgs {
  int mode = QM_CELL;
  int64_t n = 0;
  const double* r = nullptr;        // (3, n)
  const uint32_t* perm = nullptr;   // processing order (Morton-sorted), or null
  int32_t* i_cell = nullptr;        // inout (const for QM_INTERP_ICELL)
  double* out_f64 = nullptr;
  double* out2_f64 = nullptr;       // QM_INTERP_CELL: cell-data scalar per query
  int32_t* out_i32 = nullptr;
  int cell_var = 0;                 // QM_INTERP_CELL: 0-based cell_data column
  int n_vars = 0;
  int ivars[MAX_VARS] = {0};        // 0-based columns
  int max_steps = 1 << 20;
  int dynamic = 0;                  // 1: draw tiles of consecutive queries from a ticket counter (sorted batches only)
  int tiles_per_warp = 0;           // dynamic = 1: warps retire after this many tiles (0 = persistent CTAs)
  unsigned long long* ticket = nullptr;   // dynamic = 1: the ticket counter to use; null = the grid's own (every existing caller)
};

struct WalkArgs {
  int64_t n = 0;
  int max_steps = 0;
  const double* r0 = nullptr;
  const double* r1 = nullptr;
  const int32_t* ic0 = nullptr;
  const uint32_t* perm = nullptr;   // processing order (Morton order of r0), or null
  int32_t* ic1 = nullptr;
  double* r_p = nullptr;
  int32_t* n_steps = nullptr;
  int32_t* status = nullptr;
  int mask_col = -1;                // 0-based icell_data column, -1 = absent
  bool mark_early = false;          // write n_steps = INT32_MIN for rays shorter than tiny_distance (host pipeline)
};

// iu_integrate_along_field for n field lines (kernels.cu: k_trace).  Seeds are located and their
// field interpolated beforehand by a cold QM_INTERP pass (i_cell0, field0).
constexpr int TRACE_MAX_NVAR = 3;
struct TraceArgs {
  int64_t n = 0;
  int ndim = 0, nvar = 0, integrand = 0;
  int ivars[3] = {0, 0, 0};         // 0-based point_data columns of the field components
  double min_dx = 0, max_dx = 0, rtol = 0, atol = 0;
  int max_steps = 0;
  int reverse = 0, axisymmetric = 0;
  int mask_col = -1, mask_value = 0; // 0-based icell_data column, -1 = absent
  int max_walk_steps = 1 << 20;
  double* y = nullptr;              // (ndim+nvar, max_steps, n)
  double* y_field = nullptr;        // (ndim, max_steps, n) or null
  int32_t* n_steps = nullptr;
  int32_t* boundary_material = nullptr;   // or null
  const int32_t* i_cell0 = nullptr;
  const double* field0 = nullptr;   // (ndim, n)
  const uint32_t* perm = nullptr;   // processing order (Morton order of the seeds), or null
};

// every launcher returns the number of kernels it launched (for IU_B200_STAT_KERNEL_LAUNCHES)
int launch_cell_centers(const DeviceGrid& g, const double* points, const int32_t* cells, double* centers, cudaStream_t s);
int launch_build_records(const DeviceGrid& g, const double* points, const int32_t* cells, const int32_t* neighbors,
                         const double* centers, cudaStream_t s);
int launch_export_geometry(const DeviceGrid& g, double* cell_points, double* normals, double* volume, double* centers,
                           cudaStream_t s);
int launch_export_mesh(const DeviceGrid& g, double* points, int32_t* cells, cudaStream_t s);
int launch_pack_point_data(const DeviceGrid& g, const double* cols, int col0, int n_cols, cudaStream_t s);
int launch_unpack_point_data(const DeviceGrid& g, double* cols, int col0, int n_cols, cudaStream_t s);
int launch_repack_point_data(const double* old_pd, int old_stride, double* new_pd, int new_stride, int n_cols,
                             int n_points, cudaStream_t s);
size_t sort_temp_bytes(int64_t n, int end_bit);
// hint: the batch's i_cell array or null; entries >= 1 set bit 31 of the query's permutation entry (kernels.cu: PERM_HINT_BIT)
int launch_morton_sort(const DeviceGrid& g, int64_t n, const double* r, const int32_t* hint, uint32_t* keys, uint32_t* keys_alt,
                       uint32_t* idx, uint32_t* idx_alt, void* temp, size_t temp_bytes, int* end_bit_out,
                       cudaStream_t s);
int morton_end_bit(const DeviceGrid& g);
// fills lists (n_vox x VOX_K) for the grid described by v; stats[0] = voxels flagged VOX_TREE, stats[1] = sum of list lengths
int launch_build_voxels(const DeviceGrid& g, const VoxelGrid& v, int64_t n_vox, uint32_t* lists, unsigned long long* stats,
                        cudaStream_t s);
// tree_build.cu: the nearest-centre kd-tree built on the device.  centers: device (3, n_cells); fat [4 (L-1)] doubles,
// box [8 (2L-1)] floats and orig_of_dev [L LEAF] int32 are device arrays of the caller, L = ceil(n_cells / LEAF).
// Returns the number of kernels launched, or -(cudaError_t) - 1000 on a CUDA error.  Synchronises s.
int build_kdtree_device(const double* centers, int n_cells, double* fat, float* box, int32_t* orig_of_dev, cudaStream_t s);
int launch_query(const DeviceGrid& g, const QueryArgs& a, cudaStream_t s);
long long bounds_errors();   // builds with -DIU_CHECKED=1: index-range violations seen by the kernels so far; otherwise -1
int launch_walk(const DeviceGrid& g, const WalkArgs& a, cudaStream_t s);
// The sorted, ticket-scheduled query pass with ALL of its scratch (sort keys, permutation, radix-sort temporary storage,
// ticket counter) in a file-scope static __device__ array of kernels.cu: no cudaMalloc / cudaFree anywhere on the path.
// a.n <= NOALLOC_MAX_QUERIES.  The device of the grid must be current.  Returns the number of kernels launched, or a
// negative value with set_error() called.
constexpr int64_t NOALLOC_MAX_QUERIES = 1 << 20;
int launch_query_noalloc(const DeviceGrid& g, QueryArgs a, cudaStream_t s);
int launch_ray_length_sample(int64_t n, int n_samples, const double* r0, const double* r1, double* sum, cudaStream_t s);
int launch_trace_seeds(int64_t n, int ndim, int nt, int max_steps, const double* y, double* r, cudaStream_t s);
int launch_trace(const DeviceGrid& g, const TraceArgs& a, cudaStream_t s);
int launch_generate_queries(int64_t n, int64_t first, uint64_t seed, const double lo[3], const double hi[3], double* r,
                            cudaStream_t s);
int launch_merge_rows(int64_t n, int row_bytes, const int32_t* i_cell, const void* src, void* dst, cudaStream_t s);

}  // namespace iu
