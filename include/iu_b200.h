/* iu_b200.h -- C ABI of libiu_b200.so: batched point location + interpolation on unstructured
 * grids (triangle / quad / tetrahedron) on NVIDIA B200 (sm_100a).
 *
 * The reference (jannisteunissen/interpolate_unstructured) has no FFI layer: its boundary is the
 * public interface of the Fortran module m_interp_unstructured (m_interp_unstructured.f90:77-101).
 * Every entry point below names the reference procedure it replaces.  A Fortran caller binds
 * these with ISO_C_BINDING (interpolate_unstructured_b200/fortran/m_iu_b200.f90).
 *
 * Conventions (identical to the reference unless stated):
 *   - all cell / point / variable indices are 1-based; i_cell < 1 on input = "no hint" (cold
 *     start through the kd-tree), i_cell <= 0 on output = not found (left the domain);
 *   - r, r0, r1, r_p are (3, n) Fortran arrays = n consecutive (x, y, z) triples;
 *   - var_at_r is (n_vars, n) Fortran = n consecutive rows of n_vars doubles; rows (and res
 *     entries) of not-found queries are left untouched (m_interp_unstructured.f90:447, :461, :492);
 *   - point_data / cell_data / icell_data tables are column-major (n_points|n_cells, n_columns);
 *   - every function returns 0 on success and a negative IU_B200_E* code on failure with a
 *     message retrievable through iu_b200_last_error() (the reference would `error stop`);
 *   - plain functions take HOST pointers and run H2D -> kernels -> D2H internally (chunked and
 *     overlapped); *_dev functions take DEVICE pointers (on the grid's device) and enqueue on the
 *     given cudaStream_t (passed as void*; NULL = default stream) without synchronising,
 *     except where noted.
 * There is no CPU fallback: every function fails with IU_B200_ECUDA when no usable GPU exists.
 */
#ifndef IU_B200_H
#define IU_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* cell types, m_interp_unstructured.f90:13-15 */
#define IU_B200_TRIANGLE 1
#define IU_B200_QUAD 2
#define IU_B200_TETRA 3

/* error codes */
#define IU_B200_OK 0
#define IU_B200_EINVAL (-1)  /* bad argument                                              */
#define IU_B200_ECUDA (-2)   /* CUDA error / no device                                    */
#define IU_B200_EIO (-3)     /* file error ("Wrong file format", missing entry, ...)      */
#define IU_B200_EHINT (-4)   /* some i_cell > n_cells   (error stop at :490)              */
#define IU_B200_ENOMEM (-5)

/* per-query sentinel written to i_cell when the ray walk exceeded the step cap (the reference
 * walks with max_steps = huge(1), :431, and would not terminate on an invalid mesh) */
#define IU_B200_CELL_WALK_LIMIT (-2)

/* options for iu_b200_set_option */
#define IU_B200_OPT_SORT_QUERIES 1  /* 0 never, 1 always, 2 auto (default): sort batches along a space-filling curve
                                       (auto: point queries always; rays of iu_get_cell_through_neighbors in calls
                                       of >= 2^20 rays when the latest sample of ray lengths is above ~half a mean
                                       cell size - the grid's first such call waits for its sample, later ones use
                                       the latest sample that has arrived and stay asynchronous);
                                       3 = the caller's queries are already in a spatially coherent
                                       order (no sort, no permutation, ticket-tile scheduling) */
#define IU_B200_OPT_MAX_WALK_STEPS 2 /* step cap used where the reference passes huge(1); default 2^20 */
#define IU_B200_OPT_HOST_CHUNK 3    /* queries per H2D/compute/D2H pipeline chunk (host API)      */
#define IU_B200_OPT_DEV_BATCH 4     /* max queries per internal device sub-batch (workspace bound) */
#define IU_B200_OPT_TIME_KERNELS 5  /* 1: bracket the hot kernel and the sort with CUDA events (stats 7..9) */
#define IU_B200_OPT_DYNAMIC_TILES 6 /* 1 (default): sorted batches are handed to warps as tiles of consecutive
                                       queries from a device ticket counter; 0: static grid-stride assignment;
                                       2: ticket tiles with lane refill (measured slower, kept for A/B) */

#define IU_B200_OPT_NN_VOXEL_DENSITY 7 /* voxels of the nearest-centre table per 100 cells (default 400; 0 = no table,
                                       every cold query searches the kd-tree).  The table lists, per voxel of a regular
                                       grid over the mesh's bounding box, every centre that can be nearest to a point of
                                       the voxel; results are identical with and without it.  Setting it rebuilds the table. */

#define IU_B200_OPT_OVERLAP_SORT 8   /* 1 (default 0: measured, no gain): in a device batch of several sub-batches the curve keys + radix sort
                                       of sub-batch k+1 run on a second stream beside the query kernel of sub-batch k */
#define IU_B200_OPT_TILES_PER_WARP 9 /* tiles a warp of the query kernel processes before it retires while a sort runs
                                       beside it (default 8); short-lived CTAs are what lets the two kernels share SMs */

/* indices into the int64 array filled by iu_b200_get_stats */
#define IU_B200_STAT_KERNEL_LAUNCHES 0 /* kernels of this library launched since creation/reset */
#define IU_B200_STAT_WALK_LIMIT_HITS 1
#define IU_B200_STAT_QUERIES 2
#define IU_B200_STAT_WORKSPACE_BYTES 3
#define IU_B200_STAT_MESH_BYTES 4
#define IU_B200_STAT_TREE_DEPTH 5
#define IU_B200_STAT_NOT_FOUND 6       /* queries that ended with i_cell <= 0 */
#define IU_B200_STAT_QUERY_KERNEL_NS 7  /* with IU_B200_OPT_TIME_KERNELS: summed device time of the hot kernel */
#define IU_B200_STAT_QUERY_KERNEL_LAUNCHES 8
#define IU_B200_STAT_SORT_NS 9         /* same for curve keys + radix sort */
#define IU_B200_STAT_NN_VOXELS 10       /* voxels of the nearest-centre table (0 = none) */
#define IU_B200_STAT_NN_VOXELS_TREE 11  /* voxels with too many candidates: their queries search the kd-tree */
#define IU_B200_STAT_NN_VOXEL_ENTRIES 12 /* sum of the candidate-list lengths */
#define IU_B200_STAT_NN_TREE_QUERIES 13 /* cold queries that searched the kd-tree although a table exists */
#define IU_B200_STAT_NN_VOXEL_BUILD_US 14 /* device time of the last table build */
#define IU_B200_STAT_DEVICES 15        /* GPUs holding a replica of the mesh (iu_b200_device_count) */
#define IU_B200_STAT_REPLICATE_US 16   /* host time spent replicating the built mesh device to device */
#define IU_B200_STAT_BOUNDS_ERRORS 17  /* -1 in normal builds; builds with -DIU_CHECKED=1 range-check every index the kernels
                                        * derive from mesh / query data and count the violations here */
#define IU_B200_STAT_H2D_SKIPPED_BYTES 18 /* host API: bytes of i_cell NOT uploaded because a whole pipeline chunk carried no
                                           * hint (the device buffer is zeroed instead) */
#define IU_B200_STAT_TREE_BUILD_US 19    /* wall time of the kd-tree build at grid creation (device build; IU_B200_HOST_TREE=1: host) */
#define IU_B200_N_STATS 20

typedef struct iu_b200_grid iu_b200_grid; /* opaque; replaces type(iu_grid_t), :24-59 */

/* In-memory mesh description = the state iu_read_grid has after reading the file and shifting
 * indices to 1-based (m_interp_unstructured.f90:833-911).  All pointers are host pointers. */
typedef struct iu_b200_grid_desc {
  int32_t cell_type;        /* IU_B200_TRIANGLE | _QUAD | _TETRA                         */
  int32_t n_points;
  int32_t n_cells;
  const double* points;     /* (3, n_points)                                             */
  const int32_t* cells;     /* (ppc, n_cells), 1-based point indices                     */
  const int32_t* neighbors; /* (ppc, n_cells), 1-based cell across face k, < 1 = none    */
  int32_t n_point_data;
  const double* point_data; /* (n_points, n_point_data) column-major, may be NULL if 0   */
  int32_t n_cell_data;
  const double* cell_data;  /* (n_cells, n_cell_data)                                    */
  int32_t n_icell_data;
  const int32_t* icell_data; /* (n_cells, n_icell_data)                                  */
  int32_t device;           /* CUDA device ordinal, -1 = current device                  */
} iu_b200_grid_desc;

/* ---- lifecycle ---------------------------------------------------------------------------- */

/* Tail of iu_read_grid (:915-925): set_cell_points, set_face_normal_vectors, set_cell_volumes,
 * build_kdtree -- derived geometry is computed on the device with the reference's arithmetic,
 * the nearest-centre kd-tree is built and the mesh is uploaded in SoA/record form. */
int iu_b200_grid_create(const iu_b200_grid_desc* desc, iu_b200_grid** grid);

/* iu_read_grid(filename, ug, coord_scale_factor) for *.binda files (:820-927, m_binda.f90):
 * reads the BINDA container, widens float32, applies the scale factor and the +1 index shift. */
int iu_b200_grid_from_binda(const char* filename, double coord_scale_factor, int device, iu_b200_grid** grid);

/* ---- one handle, several GPUs (the mesh is replicated, queries are split, no collective) ----
 * iu_b200_grid_replicate copies the BUILT mesh (records, search structures, data tables) device to device
 * (cudaMemcpyPeer) onto every listed GPU that does not hold it yet; n_devices <= 0 or devices == NULL means every
 * visible GPU.  Afterwards every host-buffer *_batch call splits [0, n) into one contiguous slab per GPU and runs the
 * slabs concurrently (one host thread per GPU, each with its own H2D | kernels | D2H pipeline); results are identical
 * to the single-GPU handle's.  Column updates, added columns, names and options apply to every replica.  The *_dev
 * entry points take pointers of ONE device: call them with the handle of that device's replica
 * (iu_b200_grid_replica(grid, i), i = 0 .. iu_b200_device_count - 1; replica 0 is `grid` itself; info[7] of
 * iu_b200_grid_info is its device).  Replicas are owned by `grid` and destroyed with it.
 * iu_b200_grid_create_multi / _from_binda_multi = create on devices[0], then replicate. */
int iu_b200_grid_replicate(iu_b200_grid* grid, int n_devices, const int* devices);
int iu_b200_grid_create_multi(const iu_b200_grid_desc* desc, int n_devices, const int* devices, iu_b200_grid** grid);
int iu_b200_grid_from_binda_multi(const char* filename, double coord_scale_factor, int n_devices, const int* devices,
                                  iu_b200_grid** grid);
int iu_b200_device_count(const iu_b200_grid* grid);
iu_b200_grid* iu_b200_grid_replica(iu_b200_grid* grid, int i);

int iu_b200_grid_destroy(iu_b200_grid* grid);

/* message of the last failure on the calling thread */
const char* iu_b200_last_error(void);

/* info[0..7] = cell_type, n_points, n_cells, n_points_per_cell, n_point_data, n_cell_data,
 * n_icell_data, device  (iu_grid_t scalars, :25-32) */
int iu_b200_grid_info(const iu_b200_grid* grid, int32_t info[8]);
/* rmin(3), rmax(3) of the points (:867-868) */
int iu_b200_grid_bounds(const iu_b200_grid* grid, double rmin[3], double rmax[3]);

/* iu_get_point_data_index / iu_get_cell_data_index / iu_get_icell_data_index (:106-116, :224-247):
 * 1-based column of the variable called `name`, -1 if absent. */
int iu_b200_get_point_data_index(const iu_b200_grid* grid, const char* name);
int iu_b200_get_cell_data_index(const iu_b200_grid* grid, const char* name);
int iu_b200_get_icell_data_index(const iu_b200_grid* grid, const char* name);
/* Names of the columns of a grid built from arrays (iu_b200_grid_create has no name fields; iu_b200_grid_from_binda
 * takes them from the file like :894-911): kind 0 point_data, 1 cell_data, 2 icell_data; i_var 1-based. */
int iu_b200_set_data_name(iu_b200_grid* grid, int kind, int i_var, const char* name);
/* name of a column (for iu_write_vtk-style consumers); NULL if out of range */
const char* iu_b200_point_data_name(const iu_b200_grid* grid, int i_var);
const char* iu_b200_cell_data_name(const iu_b200_grid* grid, int i_var);
const char* iu_b200_icell_data_name(const iu_b200_grid* grid, int i_var);

/* iu_add_point_data / iu_add_cell_data / iu_add_icell_data (:119-161): append a zero-filled
 * column called `name`; *i_var receives its 1-based index. */
int iu_b200_add_point_data(iu_b200_grid* grid, const char* name, int* i_var);
int iu_b200_add_cell_data(iu_b200_grid* grid, const char* name, int* i_var);
int iu_b200_add_icell_data(iu_b200_grid* grid, const char* name, int* i_var);

/* Refresh device columns after the caller changed them on the host (the reference's callers
 * write ug%point_data directly, test_tetra.f90:38-40, test_trace_field.f90:20-23).
 * data = (n_points|n_cells, n_cols) column-major host array for the 1-based columns
 * col0 .. col0+n_cols-1. */
int iu_b200_update_point_data(iu_b200_grid* grid, int col0, int n_cols, const double* data);
int iu_b200_update_cell_data(iu_b200_grid* grid, int col0, int n_cols, const double* data);
int iu_b200_update_icell_data(iu_b200_grid* grid, int col0, int n_cols, const int32_t* data);

/* The reverse direction: read device columns col0 .. col0+n_cols-1 back into a column-major host
 * array (n_points|n_cells, n_cols).  The Fortran shim fills its ug%point_data / ug%cell_data /
 * ug%icell_data mirrors with these after iu_read_grid (:875-911). */
int iu_b200_download_point_data(iu_b200_grid* grid, int col0, int n_cols, double* data);
int iu_b200_download_cell_data(iu_b200_grid* grid, int col0, int n_cols, double* data);
int iu_b200_download_icell_data(iu_b200_grid* grid, int col0, int n_cols, int32_t* data);

int iu_b200_set_option(iu_b200_grid* grid, int option, int64_t value);
int iu_b200_get_stats(iu_b200_grid* grid, int64_t stats[IU_B200_N_STATS]);
int iu_b200_reset_stats(iu_b200_grid* grid);

/* ---- the hot path, host buffers ------------------------------------------------------------ */

/* find_nearby_cell_kdtree (:272-288) for n points: idx[i] = 1-based cell whose centre is nearest
 * to r(:, i) in exact fp64 (ties: lowest cell index).  Exposed for parity tests of the kd-tree. */
int iu_b200_find_nearby_cell_batch(iu_b200_grid* grid, int64_t n, const double* r, int32_t* idx);

/* iu_get_cell (:414-434) for n points; i_cell is inout. */
int iu_b200_get_cell_batch(iu_b200_grid* grid, int64_t n, const double* r, int32_t* i_cell);

/* iu_interpolate_at (:480-495) for n points.  i_vars = n_vars 1-based point_data columns. */
int iu_b200_interpolate_at_batch(iu_b200_grid* grid, int64_t n, const double* r, int n_vars, const int32_t* i_vars,
                                 double* var_at_r, int32_t* i_cell);

/* iu_interpolate_scalar_at (:465-477) for n points (n_vars = 1). */
int iu_b200_interpolate_scalar_at_batch(iu_b200_grid* grid, int64_t n, const double* r, int i_var, double* var_at_r,
                                        int32_t* i_cell);

/* iu_interpolate_at_icell (:498-527): cells already known (entries < 1 are skipped). */
int iu_b200_interpolate_at_icell_batch(iu_b200_grid* grid, int64_t n, const double* r, int n_vars,
                                       const int32_t* i_vars, double* var_at_r, const int32_t* i_cell);

/* iu_get_cell_scalar_at / iu_get_icell_scalar_at (:437-462) for n points. */
int iu_b200_get_cell_scalar_at_batch(iu_b200_grid* grid, int64_t n, const double* r, int i_var, int32_t* i_guess,
                                     double* res);
int iu_b200_get_icell_scalar_at_batch(iu_b200_grid* grid, int64_t n, const double* r, int i_var, int32_t* i_guess,
                                      int32_t* res);

/* iu_interpolate_at (:480-495) followed by iu_get_cell_scalar_at (:437-448) with the cell just found as i_guess, for n
 * points in ONE pass over the queries (the composition BASELINE config C5 times: "8 point-data vars + cell data").
 * Results are those of the two calls in sequence: var_at_r (n_vars, n) from the first call's cell; the second call
 * walks again from that cell's centre (:424-432), writes cell_res[i] = cell_data(i_cell, i_cell_var) and leaves the
 * cell it ended in in i_cell[i].  Rows of queries that were not found stay untouched.
 * (Host-buffer form only: when the second walk of a found query leaves the domain - a query within rounding of a
 * boundary face - its var_at_r row is left untouched as well; the device form writes it.) */
int iu_b200_interpolate_and_cell_scalar_at_batch(iu_b200_grid* grid, int64_t n, const double* r, int n_vars,
                                                 const int32_t* i_vars, int i_cell_var, double* var_at_r,
                                                 double* cell_res, int32_t* i_cell);

/* iu_get_cell_through_neighbors (:667-725) for n rays r0 -> r1 starting in cells ic0.
 * Outputs ic1, r_p (3, n), n_steps, status (0 reached, -1 left domain, 1 mask changed).
 * i_icell_mask = 0 means the optional argument is absent.  As in the reference, r_p and n_steps
 * of rays shorter than 1e-100 are left untouched (:689-692); after max_steps iterations without
 * termination n_steps = max_steps + 1 (Fortran do-loop semantics). */
int iu_b200_get_cell_through_neighbors_batch(iu_b200_grid* grid, int64_t n, int max_steps, const double* r0,
                                             const double* r1, const int32_t* ic0, int32_t* ic1, double* r_p,
                                             int32_t* n_steps, int32_t* status, int i_icell_mask);

/* iu_integrate_along_field (:989-1217) for n field lines: adaptive Bogacki-Shampine RK3(2) along
 * the vector field stored in point_data columns i_field[ndim] until a boundary is reached.
 *   y        (ndim+nvar, max_steps, n) inout: y(:, 1, i) holds the start position and the initial
 *            values of the nvar integrated variables; columns 2..n_steps(i) receive the curve;
 *            later columns stay untouched
 *   y_field  (ndim, max_steps, n) inout or NULL: the field along the curve
 *   n_steps  number of points stored; max_steps + 1 if the boundary was not reached (:1167-1168)
 *   i_icell_mask / mask_value: integrate only where icell_data(:, i_icell_mask) == mask_value
 *            (i_icell_mask = 0: both optional arguments absent)
 *   boundary_material (n) inout or NULL: -1 physical boundary, else the mask value of the cell
 *            behind the material boundary; untouched for lines that ran out of steps
 * The reference's callback sub_int (:61-74) cannot cross the C ABI per step on a GPU; `integrand`
 * selects a built-in one: 0 none (nvar = 0), 1 dy_var(:) = 1 (arc length, test_trace_field.f90:57-64),
 * 2 dy_var(:) = norm2(field).  ndim in 1..3, nvar in 0..3.  max_dx < min_dx and max_steps < 1 are
 * errors as in the reference (:1036-1037). */
int iu_b200_integrate_along_field_batch(iu_b200_grid* grid, int64_t n, int ndim, int nvar, int integrand,
                                        const int32_t* i_field, double min_dx, double max_dx, int max_steps,
                                        double rtol, double atol, int reverse, double* y, double* y_field,
                                        int32_t* n_steps, int axisymmetric, int i_icell_mask, int mask_value,
                                        int32_t* boundary_material);

/* ---- the hot path, device buffers (no host<->device copies; asynchronous on `stream`) ------- */

int iu_b200_find_nearby_cell_batch_dev(iu_b200_grid* grid, int64_t n, const double* r, int32_t* idx, void* stream);
int iu_b200_get_cell_batch_dev(iu_b200_grid* grid, int64_t n, const double* r, int32_t* i_cell, void* stream);
int iu_b200_interpolate_at_batch_dev(iu_b200_grid* grid, int64_t n, const double* r, int n_vars,
                                     const int32_t* i_vars /* host */, double* var_at_r, int32_t* i_cell,
                                     void* stream);
int iu_b200_interpolate_at_icell_batch_dev(iu_b200_grid* grid, int64_t n, const double* r, int n_vars,
                                           const int32_t* i_vars /* host */, double* var_at_r,
                                           const int32_t* i_cell, void* stream);
int iu_b200_interpolate_and_cell_scalar_at_batch_dev(iu_b200_grid* grid, int64_t n, const double* r, int n_vars,
                                                     const int32_t* i_vars /* host */, int i_cell_var, double* var_at_r,
                                                     double* cell_res, int32_t* i_cell, void* stream);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  iu_b200_interpolate_at_batch_dev for
 * n <= IU_B200_NOALLOC_MAX_QUERIES queries (same arguments, same results, queries processed in curve order through a
 * permutation) that allocates NOTHING: its scratch - sort keys, permutation, radix-sort temporary storage, the ticket
 * counter of the tile scheduler - lives in one static __device__ array of 24 MiB + 256 B inside the library, which every
 * GPU that loads the library holds until exit.  Larger n is rejected with IU_B200_EINVAL.  Every call shares that array
 * (whatever the grid): calls that overlap on the GPU - different streams, or two host threads with two grids - race on it;
 * calls on one stream, or otherwise ordered on the device, are fine.  The per-kernel timing option is ignored. */
#define IU_B200_NOALLOC_MAX_QUERIES 1048576
int iu_b200_interpolate_at_batch_dev_noalloc(iu_b200_grid* grid, int64_t n, const double* r, int n_vars,
                                             const int32_t* i_vars /* host */, double* var_at_r, int32_t* i_cell,
                                             void* stream);
int iu_b200_get_cell_scalar_at_batch_dev(iu_b200_grid* grid, int64_t n, const double* r, int i_var,
                                         int32_t* i_guess, double* res, void* stream);
int iu_b200_get_icell_scalar_at_batch_dev(iu_b200_grid* grid, int64_t n, const double* r, int i_var,
                                          int32_t* i_guess, int32_t* res, void* stream);
int iu_b200_get_cell_through_neighbors_batch_dev(iu_b200_grid* grid, int64_t n, int max_steps, const double* r0,
                                                 const double* r1, const int32_t* ic0, int32_t* ic1, double* r_p,
                                                 int32_t* n_steps, int32_t* status, int i_icell_mask,
                                                 void* stream);
int iu_b200_integrate_along_field_batch_dev(iu_b200_grid* grid, int64_t n, int ndim, int nvar, int integrand,
                                            const int32_t* i_field /* host */, double min_dx, double max_dx,
                                            int max_steps, double rtol, double atol, int reverse, double* y,
                                            double* y_field, int32_t* n_steps, int axisymmetric, int i_icell_mask,
                                            int mask_value, int32_t* boundary_material, void* stream);
/* The *_dev calls cannot return IU_B200_EHINT synchronously.  This call synchronises `stream`
 * and returns IU_B200_EHINT if any query since the last check carried i_cell > n_cells. */
int iu_b200_check_dev(iu_b200_grid* grid, void* stream);

/* ---- utilities ------------------------------------------------------------------------------ */

/* Face-neighbour table as convert_to_binary.py:118-162 computes it (coincident points merged,
 * face i = vertices (i+k) mod ppc, neighbour only if exactly two cells share the face).
 * cells and neighbors are 0-based here (on-disk convention), -1 = no neighbour. */
int iu_b200_build_cell_neighbors(int cell_type, int32_t n_points, int32_t n_cells, const double* points,
                                 const int32_t* cells, int32_t* neighbors);

/* Derived geometry in ORIGINAL cell order, for parity checks against the reference arrays
 * cell_points(3,ppc,Nc), cell_face_normals(3,ppc,Nc), cell_volume(Nc), cell centres (3,Nc).
 * Any pointer may be NULL. */
int iu_b200_export_geometry(iu_b200_grid* grid, double* cell_points, double* cell_face_normals, double* cell_volume,
                            double* cell_centers);

/* The raw mesh back from the device records, ORIGINAL order: points (3, Np), cells (ppc, Nc)
 * 1-based.  Either pointer may be NULL.  points is the array the grid was created from (ug%points), so points
 * that no cell refers to are kept. */
int iu_b200_export_mesh(iu_b200_grid* grid, double* points, int32_t* cells);

/* iu_write_vtk(ug, filename) (m_interp_unstructured.f90:930-985 through m_vtk.f90:45-221): the
 * grid with its current cell, integer cell and point data as a .vtu file, byte for byte the
 * reference's stream (XML header + appended raw blocks, Float32 points, 0-based connectivity).
 * Columns without a name are called cell_data_<i> / icell_data_<i> / point_data_<i>. */
int iu_b200_write_vtk(iu_b200_grid* grid, const char* filename);
/* The same writer on host arrays (no device needed): points (3, Np), cells (ppc, Nc) 1-based,
 * data tables column-major; name arrays may be NULL. */
int iu_b200_write_vtk_arrays(const char* filename, int cell_type, int32_t n_points, int32_t n_cells, const double* points,
                             const int32_t* cells, int n_cell_data, const double* cell_data, const char* const* cell_names,
                             int n_icell_data, const int32_t* icell_data, const char* const* icell_names,
                             int n_point_data, const double* point_data, const char* const* point_names);

/* n query points uniform in [lo, hi] from a counter-based generator (SplitMix64 of
 * 3*(first+i)+k, seed), written to device memory r (3, n); used by the benchmark for the
 * configurations whose query sets are too large to cross PCIe. */
int iu_b200_generate_queries_dev(iu_b200_grid* grid, int64_t n, int64_t first, uint64_t seed, const double lo[3],
                                 const double hi[3], double* r, void* stream);

/* page-locked host memory for the host-buffer API (plain malloc'd memory also works, slower) */
void* iu_b200_host_alloc(size_t bytes);
void iu_b200_host_free(void* p);

int iu_b200_version(void);

#ifdef __cplusplus
}
#endif
#endif /* IU_B200_H */
