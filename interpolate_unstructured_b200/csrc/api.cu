// api.cu -- the C ABI of libiu_b200.so (include/iu_b200.h): grid lifecycle, the device-buffer
// entry points (sub-batching + optional Morton sort around the kernels of kernels.cu) and the
// host-buffer entry points (three-slot H2D / compute / D2H pipeline on three streams).
//
// There is no CPU fallback anywhere in this file: every entry point needs a CUDA device.
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <mutex>
#include <new>
#include <thread>

#include "../../include/iu_b200.h"
#include "device.h"

namespace iu {

static thread_local std::string g_error;
void set_error(const std::string& msg) { g_error = msg; }

namespace {

struct CudaFail {
  int code;
};

#define CU(expr)                                                                                         \
  do {                                                                                                   \
    cudaError_t e__ = (expr);                                                                            \
    if (e__ != cudaSuccess) {                                                                            \
      set_error(std::string(#expr) + ": " + cudaGetErrorString(e__));                                    \
      throw CudaFail{e__ == cudaErrorMemoryAllocation ? IU_B200_ENOMEM : IU_B200_ECUDA};                 \
    }                                                                                                    \
  } while (0)

template <class T>
struct DevBuf {   // RAII device buffer
  T* p = nullptr;
  size_t n = 0;
  DevBuf() = default;
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  ~DevBuf() { release(); }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    n = 0;
  }
  void alloc(size_t count) {
    release();
    if (count == 0) return;
    CU(cudaMalloc((void**)&p, count * sizeof(T)));
    n = count;
  }
  void ensure(size_t count) {
    if (count > n) alloc(count);
  }
  T* take() {
    T* q = p;
    p = nullptr;
    n = 0;
    return q;
  }
};

constexpr int N_SLOTS = 3;   // host-API pipeline depth

struct Slot {
  DevBuf<double> r, r1, out_f64;
  DevBuf<int32_t> i_a, i_b, i_c, i_d;
  cudaEvent_t in_done = nullptr, comp_done = nullptr, out_done = nullptr;
};

}  // namespace
}  // namespace iu

using namespace iu;

struct iu_b200_grid {
  int device = 0;
  DeviceGrid dg;
  std::vector<std::string> point_names, cell_names, icell_names;
  // options
  int sort_mode = 2;
  int64_t max_walk_steps = 1 << 20;
  int64_t host_chunk = 1 << 22;
  int64_t dev_batch = 1 << 26;
  int dynamic_tiles = 1;   // sorted sub-batches draw tiles of queries from a ticket counter (kernels.cu, SCHED = 1)
  int voxel_density = 400; // voxels of the nearest-centre table per 100 cells (0 = no table, kd-tree search only)
  double cmin[3] = {0, 0, 0}, cmax[3] = {0, 0, 0};   // bounds of the cell centres
  int64_t vox_flagged = 0, vox_entries = 0;
  double vox_build_ms = 0;
  // stats
  int64_t n_launches = 0, n_queries = 0;
  size_t mesh_bytes = 0;
  int tree_depth = 0;
  // sort workspace (device API and host pipeline share it: all compute is stream-ordered by the caller)
  DevBuf<uint32_t> keys, keys_alt, idx, idx_alt;
  DevBuf<char> sort_temp;
  // sort of sub-batch k+1 beside k_query of sub-batch k (run_query): second permutation buffer, a high-priority stream
  // for the sorts and the events that order the two streams
  DevBuf<uint32_t> idx_alt2;
  cudaStream_t s_sort = nullptr;
  cudaEvent_t ev_enter = nullptr, ev_sorted[2] = {nullptr, nullptr}, ev_queried[2] = {nullptr, nullptr};
  size_t h2d_skipped_bytes = 0;   // i_cell uploads replaced by a device memset (all-cold chunks of the host API)
  int overlap_sort = 0;        // IU_B200_OPT_OVERLAP_SORT (measured: no gain, profiles/experiments/r02k_sort_overlap.txt)
  int tiles_per_warp = 8;      // k_query warps retire after this many tiles while a sort runs beside them
  // ray-walk sort decision (IU_B200_OPT_SORT_QUERIES = 2): mean sampled ray length of the latest measured call
  DevBuf<double> walk_sum;
  double* h_walk_sum = nullptr;   // pinned
  cudaEvent_t walk_ev = nullptr;
  bool walk_ev_pending = false, walk_measured = false, walk_sort = false;
  // tracer workspace: seeds (3, n), their field (ndim, n) and cells
  DevBuf<double> tr_seed, tr_field;
  DevBuf<int32_t> tr_cell;
  // host pipeline
  cudaStream_t s_in = nullptr, s_comp = nullptr, s_out = nullptr;
  Slot slot[N_SLOTS];
  unsigned long long* h_counters = nullptr;   // pinned, N_SLOTS x CNT_N snapshots
  void* h_stage = nullptr;                     // pinned staging for the rare "merge" path
  size_t h_stage_bytes = 0;
  unsigned long long seen[CNT_N] = {0};   // counter values already accounted for
  // optional per-kernel timing (IU_B200_OPT_TIME_KERNELS): event pairs around the hot kernel and the sort
  bool time_kernels = false;
  struct Timed { cudaEvent_t a, b; int kind; };
  std::vector<Timed> pending;
  std::vector<cudaEvent_t> event_pool;
  double timed_ns[2] = {0, 0};
  int64_t timed_launches[2] = {0, 0};
  std::mutex mu;   // one batch call at a time per grid (workspaces are per grid)
  // The sort buffers, the ticket counter and the tracer workspace are per grid: a call enqueued on another stream
  // than the previous one first waits (on the device) for that call's last kernel.
  cudaEvent_t ws_done = nullptr;
  cudaStream_t ws_stream = nullptr;
  bool ws_used = false;
  // Further replicas of the mesh on other GPUs (iu_b200_grid_replicate): full single-device grids of their own.  The
  // host-buffer batch calls split [0, n) into one contiguous slab per device and run them concurrently.
  std::vector<iu_b200_grid*> replicas;
  double replicate_ms = 0, tree_build_ms = 0;
};

namespace {

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
    else prev = -1;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

size_t workspace_bytes(const iu_b200_grid* g) {
  size_t b = (g->keys.n + g->keys_alt.n + g->idx.n + g->idx_alt.n + g->idx_alt2.n) * 4 + g->sort_temp.n + g->h_stage_bytes;
  for (const Slot& s : g->slot)
    b += (s.r.n + s.r1.n + s.out_f64.n) * 8 + (s.i_a.n + s.i_b.n + s.i_c.n + s.i_d.n) * 4;
  return b;
}

cudaEvent_t timing_event(iu_b200_grid* g) {
  cudaEvent_t e = nullptr;
  if (!g->event_pool.empty()) { e = g->event_pool.back(); g->event_pool.pop_back(); }
  else CU(cudaEventCreate(&e));
  return e;
}

// Brackets every use of the per-grid device workspace on stream s (run_query / run_walk / run_trace).
struct WorkspaceScope {
  iu_b200_grid* g; cudaStream_t s;
  WorkspaceScope(iu_b200_grid* g_, cudaStream_t s_) : g(g_), s(s_) {
    if (!g->ws_done) CU(cudaEventCreateWithFlags(&g->ws_done, cudaEventDisableTiming));
    if (g->ws_used && g->ws_stream != s) CU(cudaStreamWaitEvent(s, g->ws_done, 0));
  }
  ~WorkspaceScope() {
    cudaEventRecord(g->ws_done, s);
    g->ws_stream = s;
    g->ws_used = true;
  }
};

// kind 0 = hot query kernel, 1 = Morton keys + radix sort
struct TimedScope {
  iu_b200_grid* g; cudaStream_t s; int kind; cudaEvent_t a = nullptr;
  TimedScope(iu_b200_grid* g_, cudaStream_t s_, int kind_) : g(g_), s(s_), kind(kind_) {
    if (g->time_kernels) { a = timing_event(g); CU(cudaEventRecord(a, s)); }
  }
  void stop() {
    if (!a) return;
    cudaEvent_t b = timing_event(g);
    CU(cudaEventRecord(b, s));
    g->pending.push_back({a, b, kind});
    a = nullptr;
  }
};

void collect_timings(iu_b200_grid* g) {
  for (auto& t : g->pending) {
    float ms = 0;
    if (cudaEventSynchronize(t.b) == cudaSuccess && cudaEventElapsedTime(&ms, t.a, t.b) == cudaSuccess) {
      g->timed_ns[t.kind] += (double)ms * 1e6;
      g->timed_launches[t.kind] += 1;
    }
    g->event_pool.push_back(t.a);
    g->event_pool.push_back(t.b);
  }
  g->pending.clear();
}

void free_device_grid(DeviceGrid& d) {
  cudaFree(d.cen32); cudaFree(d.vox_lists); cudaFree(d.points);
  cudaFree(d.fat); cudaFree(d.box); cudaFree(d.centers4); cudaFree(d.rec); cudaFree(d.dev_of_orig);
  cudaFree(d.orig_of_dev); cudaFree(d.pd); cudaFree(d.cd); cudaFree(d.icd); cudaFree(d.counters);
  d = DeviceGrid();
}

int round_up_stride(int n_cols) {
  // point-major rows: keep rows 16-byte aligned and, from 3 columns on, 32-byte (sector) aligned
  if (n_cols <= 1) return 1;
  if (n_cols == 2) return 2;
  return (n_cols + 3) / 4 * 4;
}

template <class F>
int guarded(F&& f) {
  try {
    return f();
  } catch (const CudaFail& e) {
    return e.code;
  } catch (const std::bad_alloc&) {
    set_error("out of host memory");
    return IU_B200_ENOMEM;
  }
}

// applies a mutation to the further replicas of a multi-GPU handle once it succeeded on the primary
template <class F>
int on_replicas(iu_b200_grid* g, int rc, F&& f) {
  if (rc != IU_B200_OK) return rc;
  for (iu_b200_grid* r : g->replicas) {
    const int rc2 = f(r);
    if (rc2 != IU_B200_OK) return rc2;
  }
  return IU_B200_OK;
}

// Host-buffer batch calls of a multi-GPU handle: [0, n) is cut into one contiguous slab per device (multiples of 4
// queries, so that every slab keeps the 32-byte alignment of the result rows) and f(grid of that device, offset, count)
// runs on one host thread per device.  There is no exchange between the devices: every replica holds the whole mesh.
constexpr int64_t MULTI_MIN_QUERIES = 1 << 16;
template <class F>
int run_slabs(iu_b200_grid* g, int64_t n, F&& f, int64_t min_n = MULTI_MIN_QUERIES) {
  const int nd = g ? 1 + (int)g->replicas.size() : 1;
  if (nd == 1 || n < min_n) return f(g, (int64_t)0, n);
  const int64_t slab = ((n + nd - 1) / nd + 3) / 4 * 4;
  std::vector<int> rcs(nd, IU_B200_OK);
  std::vector<std::string> errs(nd);
  std::vector<std::thread> threads;
  for (int d = 0; d < nd; ++d) {
    const int64_t off = (int64_t)d * slab, m = std::min(slab, n - off);
    if (m <= 0) break;
    iu_b200_grid* gd = d == 0 ? g : g->replicas[d - 1];
    threads.emplace_back([&, d, gd, off, m]() {
      rcs[d] = f(gd, off, m);
      if (rcs[d] != IU_B200_OK) errs[d] = g_error;
    });
  }
  for (std::thread& t : threads) t.join();
  for (int d = 0; d < nd; ++d)
    if (rcs[d] != IU_B200_OK) { set_error(errs[d]); return rcs[d]; }
  return IU_B200_OK;
}

bool check_device_available() {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    set_error(std::string("no usable CUDA device (") + (e != cudaSuccess ? cudaGetErrorString(e) : "count = 0") +
              "); libiu_b200 has no CPU fallback");
    return false;
  }
  return true;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// lifecycle
// ------------------------------------------------------------------------------------------------
namespace {

// (Re)builds the voxel table of the nearest-centre search for g->voxel_density voxels per 100 cells (device.h:
// VoxelGrid, kernels.cu: k_build_voxels).  The table is an accelerator only: without it (density 0, degenerate
// boxes, not enough memory) every cold query runs the exact kd-tree search.
void build_voxel_table(iu_b200_grid* g) {
  DeviceGrid& d = g->dg;
  if (d.vox_lists) { CU(cudaDeviceSynchronize()); cudaFree(d.vox_lists); }
  d.vox_lists = nullptr;
  d.vox = VoxelGrid();
  d.n_vox = 0;
  g->vox_flagged = g->vox_entries = 0;
  g->vox_build_ms = 0;
  if (g->voxel_density <= 0 || d.n_cells < 1) return;
  VoxelGrid v;
  int nd = 0;
  double vol = 1.0;
  for (int k = 0; k < 3; ++k) {
    const double e = d.rmax[k] - d.rmin[k];
    if (!(e >= 0.0) || !std::isfinite(e)) return;
    v.lo[k] = d.rmin[k];
    if (e == 0.0) {
      if (g->cmin[k] != g->cmax[k]) return;   // cannot happen for points that share the coordinate; stay exact anyway
      v.flat_mask |= 1 << k;
      v.flat_val[k] = g->cmin[k];
    } else { vol *= e; ++nd; }
  }
  if (nd == 0 || !(vol > 0.0) || !std::isfinite(vol)) return;
  size_t free_b = 0, total_b = 0;
  CU(cudaMemGetInfo(&free_b, &total_b));
  double want = (double)d.n_cells * g->voxel_density / 100.0;
  const double cap = std::min((double)free_b / 4.0, (double)total_b / 8.0) / (VOX_K * 4.0);
  want = std::max(1.0, std::min(want, std::min(cap, 1.5e9)));
  const double hv = std::pow(vol / want, 1.0 / nd);
  int64_t n_vox = 1;
  for (int k = 0; k < 3; ++k) {
    if (v.flat_mask >> k & 1) { v.n[k] = 1; v.inv[k] = 0.0; v.h[k] = 0.0; continue; }
    const double e = d.rmax[k] - d.rmin[k];
    v.n[k] = (int)std::max(1.0, std::min(4096.0, std::ceil(e / hv)));
    v.h[k] = e / v.n[k];
    v.inv[k] = v.n[k] / e;
    // the inflation of the voxel boxes (1e-3 h) must dwarf the rounding of lo + i*h at these coordinates
    const double mag = std::max(std::fabs(d.rmin[k]), std::fabs(d.rmax[k]));
    if (!(v.h[k] * 1e-3 > 1e4 * mag * 2.3e-16) || !std::isfinite(v.inv[k])) return;
    n_vox *= v.n[k];
  }
  if ((double)n_vox > cap) return;
  DevBuf<uint32_t> lists;
  DevBuf<unsigned long long> st;
  lists.alloc((size_t)n_vox * VOX_K);
  st.alloc(2);
  CU(cudaMemset(st.p, 0, 2 * sizeof(unsigned long long)));
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
  CU(cudaEventRecord(e0, nullptr));
  g->n_launches += launch_build_voxels(d, v, n_vox, lists.p, st.p, nullptr);
  CU(cudaEventRecord(e1, nullptr));
  CU(cudaGetLastError());
  CU(cudaDeviceSynchronize());
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  unsigned long long h[2];
  CU(cudaMemcpy(h, st.p, sizeof(h), cudaMemcpyDeviceToHost));
  g->vox_flagged = (int64_t)h[0];
  g->vox_entries = (int64_t)h[1];
  g->vox_build_ms = ms;
  d.n_vox = n_vox;
  d.vox_lists = lists.take();
  v.lists = d.vox_lists;
  d.vox = v;
}

// streams, events and pinned counter snapshots of the host-buffer pipeline (current device = g->device)
void init_pipeline(iu_b200_grid* g) {
  CU(cudaStreamCreateWithFlags(&g->s_in, cudaStreamNonBlocking));
  CU(cudaStreamCreateWithFlags(&g->s_comp, cudaStreamNonBlocking));
  CU(cudaStreamCreateWithFlags(&g->s_out, cudaStreamNonBlocking));
  for (Slot& sl : g->slot) {
    CU(cudaEventCreateWithFlags(&sl.in_done, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&sl.comp_done, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&sl.out_done, cudaEventDisableTiming));
  }
  CU(cudaMallocHost((void**)&g->h_counters, N_SLOTS * CNT_N * sizeof(unsigned long long)));
}

// A replica of a built grid on another GPU: every device array is copied device to device (cudaMemcpyPeer: over
// NVLink where the GPUs are peers), nothing is rebuilt.  The replica is a complete single-device grid.
iu_b200_grid* clone_to_device(iu_b200_grid* src, int device) {
  iu_b200_grid* g = new iu_b200_grid();
  g->device = device;
  try {
    DeviceGuard guard(device);
    // direct peer copies (NVLink / NVSwitch): without peer access cudaMemcpyPeer stages through host memory
    // (measured 28 GB/s for the 30.7 GB of the C5 mesh)
    if (device != src->device) {
      int can = 0;
      if (cudaDeviceCanAccessPeer(&can, device, src->device) == cudaSuccess && can) {
        const cudaError_t e = cudaDeviceEnablePeerAccess(src->device, 0);
        if (e != cudaSuccess) cudaGetLastError();   // already enabled (or refused: the staged copy still works)
      }
    }
    const DeviceGrid& a = src->dg;
    DeviceGrid& d = g->dg;
    d = a;                                   // scalars; the pointers are replaced below
    d.fat = d.centers4 = d.cen32 = d.rec = d.points = d.pd = d.cd = nullptr;
    d.vox_lists = nullptr; d.vox.lists = nullptr; d.box = nullptr;
    d.dev_of_orig = d.orig_of_dev = nullptr; d.icd = nullptr; d.counters = nullptr;
    auto copy = [&](auto*& dst, const auto* from, size_t bytes) {
      if (!from || bytes == 0) return;
      CU(cudaMalloc((void**)&dst, bytes));
      CU(cudaMemcpyPeerAsync(dst, device, from, src->device, bytes, nullptr));
    };
    const size_t n_slots = (size_t)a.n_leaves * LEAF;
    copy(d.fat, a.fat, (size_t)a.n_internal * 32);
    copy(d.centers4, a.centers4, n_slots * 24);
    copy(d.cen32, a.cen32, n_slots * 32);
    copy(d.box, a.box, std::max<size_t>(32, ((size_t)a.n_internal + a.n_leaves) * 8 * sizeof(float)));
    copy(d.rec, a.rec, n_slots * rec_stride(a.ppc) * 8);
    copy(d.points, a.points, (size_t)a.n_points * 24);
    copy(d.dev_of_orig, a.dev_of_orig, ((size_t)a.n_cells + 1) * 4);
    copy(d.orig_of_dev, a.orig_of_dev, n_slots * 4);
    copy(d.vox_lists, a.vox_lists, (size_t)a.n_vox * VOX_K * 4);
    d.vox.lists = d.vox_lists;
    copy(d.pd, a.pd, (size_t)a.n_points * a.pd_stride * 8);
    copy(d.cd, a.cd, (size_t)a.n_cells * a.cd_cap * 8);
    copy(d.icd, a.icd, (size_t)a.n_cells * a.icd_cap * 4);
    CU(cudaMalloc((void**)&d.counters, CNT_N * sizeof(unsigned long long)));
    CU(cudaMemsetAsync(d.counters, 0, CNT_N * sizeof(unsigned long long), nullptr));
    CU(cudaDeviceSynchronize());
    g->point_names = src->point_names; g->cell_names = src->cell_names; g->icell_names = src->icell_names;
    g->sort_mode = src->sort_mode; g->max_walk_steps = src->max_walk_steps; g->host_chunk = src->host_chunk;
    g->dev_batch = src->dev_batch; g->dynamic_tiles = src->dynamic_tiles; g->voxel_density = src->voxel_density;
    g->time_kernels = src->time_kernels; g->overlap_sort = src->overlap_sort; g->tiles_per_warp = src->tiles_per_warp;
    for (int k = 0; k < 3; ++k) { g->cmin[k] = src->cmin[k]; g->cmax[k] = src->cmax[k]; }
    g->vox_flagged = src->vox_flagged; g->vox_entries = src->vox_entries; g->vox_build_ms = 0;
    g->mesh_bytes = src->mesh_bytes; g->tree_depth = src->tree_depth;
    init_pipeline(g);
  } catch (...) {
    iu_b200_grid_destroy(g);
    throw;
  }
  return g;
}

int create_from_arrays(int cell_type, int n_points, int n_cells, const double* points, const int32_t* cells,
                       const int32_t* neighbors, int n_pd, const double* pd, int n_cd, const double* cd, int n_icd,
                       const int32_t* icd, int device, iu_b200_grid** out) {
  if (!out) { set_error("grid pointer is null"); return IU_B200_EINVAL; }
  *out = nullptr;
  if (cell_type < 1 || cell_type > 3) { set_error("unknown cell type"); return IU_B200_EINVAL; }
  if (n_points < 1 || n_cells < 1 || !points || !cells || !neighbors) { set_error("empty or missing mesh arrays"); return IU_B200_EINVAL; }
  if (n_pd < 0 || n_cd < 0 || n_icd < 0 || (n_pd > 0 && !pd) || (n_cd > 0 && !cd) || (n_icd > 0 && !icd)) {
    set_error("inconsistent data tables");
    return IU_B200_EINVAL;
  }
  const int ppc = cell_type == IU_TRIANGLE_CT ? 3 : 4;
  for (size_t i = 0; i < (size_t)n_cells * ppc; ++i) {
    if (cells[i] < 1 || cells[i] > n_points) { set_error("cells: point index outside 1..n_points"); return IU_B200_EINVAL; }
    if (neighbors[i] > n_cells) { set_error("neighbors: cell index above n_cells"); return IU_B200_EINVAL; }
  }
  if (!check_device_available()) return IU_B200_ECUDA;
  if (device < 0) cudaGetDevice(&device);
  iu_b200_grid* g = new iu_b200_grid();
  g->device = device;
  int rc = guarded([&]() {
    DeviceGuard guard(device);
    cudaStream_t s = nullptr;
    DeviceGrid& d = g->dg;
    d.cell_type = cell_type; d.ppc = ppc; d.n_cells = n_cells; d.n_points = n_points;
    for (int k = 0; k < 3; ++k) { d.rmin[k] = points[k]; d.rmax[k] = points[k]; }   // :867-868
    for (size_t i = 1; i < (size_t)n_points; ++i)
      for (int k = 0; k < 3; ++k) {
        const double v = points[3 * i + k];
        d.rmin[k] = v < d.rmin[k] ? v : d.rmin[k];
        d.rmax[k] = v > d.rmax[k] ? v : d.rmax[k];
      }
    // raw mesh -> device (temporary; only the records survive)
    DevBuf<double> d_points, d_centers;
    DevBuf<int32_t> d_cells, d_neigh;
    d_points.alloc((size_t)3 * n_points);
    d_cells.alloc((size_t)ppc * n_cells);
    d_neigh.alloc((size_t)ppc * n_cells);
    d_centers.alloc((size_t)3 * n_cells);
    CU(cudaMemcpyAsync(d_points.p, points, d_points.n * 8, cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(d_cells.p, cells, d_cells.n * 4, cudaMemcpyHostToDevice, s));
    CU(cudaMemcpyAsync(d_neigh.p, neighbors, d_neigh.n * 4, cudaMemcpyHostToDevice, s));
    // cell centres with the reference's arithmetic (build_kdtree :251-262), then the tree on the host
    g->n_launches += launch_cell_centers(d, d_points.p, d_cells.p, d_centers.p, s);
    std::vector<double> h_centers((size_t)3 * n_cells);
    CU(cudaMemcpyAsync(h_centers.data(), d_centers.p, h_centers.size() * 8, cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    for (int k = 0; k < 3; ++k) g->cmin[k] = g->cmax[k] = h_centers[k];
    for (size_t i = 1; i < (size_t)n_cells; ++i)
      for (int k = 0; k < 3; ++k) {
        const double v = h_centers[3 * i + k];
        g->cmin[k] = v < g->cmin[k] ? v : g->cmin[k];
        g->cmax[k] = v > g->cmax[k] ? v : g->cmax[k];
      }
    std::vector<double>().swap(h_centers);
    // the nearest-centre kd-tree (build_kdtree, :251-262): built on the device from the device copy of the centres
    // (tree_build.cu); IU_B200_HOST_TREE=1 selects the host builder of host_mesh.cpp (same kind of tree) for A/B
    d.n_leaves = (n_cells + LEAF - 1) / LEAF;
    d.n_internal = d.n_leaves - 1;
    for (int l = 1; l < d.n_leaves; l <<= 1) ++g->tree_depth;
    d.depth = g->tree_depth;
    const size_t n_slots = (size_t)d.n_leaves * LEAF;
    const size_t box_floats = 8 * (size_t)(2 * d.n_leaves - 1);
    CU(cudaMalloc((void**)&d.orig_of_dev, n_slots * 4));
    CU(cudaMalloc((void**)&d.dev_of_orig, ((size_t)n_cells + 1) * 4));
    CU(cudaMalloc((void**)&d.centers4, n_slots * 24));
    CU(cudaMalloc((void**)&d.cen32, n_slots * 32));
    CU(cudaMalloc((void**)&d.rec, n_slots * rec_stride(ppc) * 8));
    CU(cudaMalloc((void**)&d.counters, CNT_N * sizeof(unsigned long long)));
    CU(cudaMemsetAsync(d.counters, 0, CNT_N * sizeof(unsigned long long), s));
    CU(cudaMemsetAsync(d.dev_of_orig, 0, ((size_t)n_cells + 1) * 4, s));
    if (d.n_internal > 0) CU(cudaMalloc((void**)&d.fat, (size_t)d.n_internal * 32));
    CU(cudaMalloc((void**)&d.box, std::max<size_t>(32, box_floats * sizeof(float))));
    const auto t_tree = std::chrono::steady_clock::now();
    if (std::getenv("IU_B200_HOST_TREE")) {
      std::vector<double> hc((size_t)3 * n_cells);
      CU(cudaMemcpy(hc.data(), d_centers.p, hc.size() * 8, cudaMemcpyDeviceToHost));
      KdTreeHost tree;
      build_kdtree(hc.data(), n_cells, tree);
      if (d.n_internal > 0) {
        // nodes with their children's splits alongside: {split, split(2i+1), split(2i+2), dimensions}
        std::vector<double> fat((size_t)4 * d.n_internal, 0.0);
        for (int i = 0; i < d.n_internal; ++i) {
          unsigned long long dims = tree.sdim[i];
          fat[4 * (size_t)i] = tree.split[i];
          for (int c = 0; c < 2; ++c) {
            const long long ch = 2LL * i + 1 + c;
            if (ch < d.n_internal) { fat[4 * (size_t)i + 1 + c] = tree.split[ch]; dims |= (unsigned long long)tree.sdim[ch] << (2 + 2 * c); }
          }
          std::memcpy(&fat[4 * (size_t)i + 3], &dims, 8);
        }
        CU(cudaMemcpy(d.fat, fat.data(), fat.size() * sizeof(double), cudaMemcpyHostToDevice));
      }
      if (!tree.box.empty()) CU(cudaMemcpy(d.box, tree.box.data(), tree.box.size() * sizeof(float), cudaMemcpyHostToDevice));
      CU(cudaMemcpy(d.orig_of_dev, tree.orig_of_slot.data(), n_slots * 4, cudaMemcpyHostToDevice));
    } else {
      const int rc_tree = build_kdtree_device(d_centers.p, n_cells, d.fat, d.box, d.orig_of_dev, s);
      if (rc_tree < 0) {
        const cudaError_t ce = (cudaError_t)(-(rc_tree + 1000));
        set_error(std::string("kd-tree build: ") + cudaGetErrorString(ce));
        throw CudaFail{ce == cudaErrorMemoryAllocation ? IU_B200_ENOMEM : IU_B200_ECUDA};
      }
      g->n_launches += rc_tree;
    }
    g->tree_build_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_tree).count();
    g->n_launches += launch_build_records(d, d_points.p, d_cells.p, d_neigh.p, d_centers.p, s);
    CU(cudaStreamSynchronize(s));
    d.points = d_points.take();
    d_cells.release(); d_neigh.release(); d_centers.release();
    if (const char* env = std::getenv("IU_B200_NN_VOXEL_DENSITY")) g->voxel_density = std::max(0, std::atoi(env));
    if (const char* env = std::getenv("IU_B200_DYNAMIC_TILES")) g->dynamic_tiles = std::min(2, std::max(0, std::atoi(env)));   // A/B knob
    if (const char* env = std::getenv("IU_B200_OVERLAP_SORT")) g->overlap_sort = std::atoi(env) != 0;                          // A/B knob
    if (const char* env = std::getenv("IU_B200_TILES_PER_WARP")) g->tiles_per_warp = std::max(1, std::atoi(env));               // A/B knob
    if (const char* env = std::getenv("IU_B200_L2_FETCH")) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)std::atoi(env));   // experiment knob
    build_voxel_table(g);
    // data tables
    d.n_pd = n_pd;
    d.pd_stride = round_up_stride(n_pd);
    if (n_pd > 0) {
      CU(cudaMalloc((void**)&d.pd, (size_t)n_points * d.pd_stride * 8));
      CU(cudaMemsetAsync(d.pd, 0, (size_t)n_points * d.pd_stride * 8, s));
      DevBuf<double> cols;
      cols.alloc((size_t)n_points * n_pd);
      CU(cudaMemcpyAsync(cols.p, pd, cols.n * 8, cudaMemcpyHostToDevice, s));
      g->n_launches += launch_pack_point_data(d, cols.p, 0, n_pd, s);
      CU(cudaStreamSynchronize(s));
    }
    d.n_cd = d.cd_cap = n_cd;
    if (n_cd > 0) {
      CU(cudaMalloc((void**)&d.cd, (size_t)n_cells * n_cd * 8));
      CU(cudaMemcpy(d.cd, cd, (size_t)n_cells * n_cd * 8, cudaMemcpyHostToDevice));
    }
    d.n_icd = d.icd_cap = n_icd;
    if (n_icd > 0) {
      CU(cudaMalloc((void**)&d.icd, (size_t)n_cells * n_icd * 4));
      CU(cudaMemcpy(d.icd, icd, (size_t)n_cells * n_icd * 4, cudaMemcpyHostToDevice));
    }
    g->mesh_bytes = (size_t)n_points * 24 + n_slots * (rec_stride(ppc) * 8 + 24 + 32 + 4) + ((size_t)n_cells + 1) * 4 + (size_t)d.n_internal * 32 + box_floats * sizeof(float) +
                    (size_t)n_points * d.pd_stride * 8 + (size_t)n_cells * (n_cd * 8 + n_icd * 4);
    init_pipeline(g);
    return IU_B200_OK;
  });
  if (rc != IU_B200_OK) {
    iu_b200_grid_destroy(g);
    return rc;
  }
  g->point_names.resize(n_pd);
  g->cell_names.resize(n_cd);
  g->icell_names.resize(n_icd);
  *out = g;
  return IU_B200_OK;
}

}  // namespace

extern "C" {

int iu_b200_version(void) { return 100; }

const char* iu_b200_last_error(void) { return g_error.c_str(); }

int iu_b200_grid_create(const iu_b200_grid_desc* desc, iu_b200_grid** grid) {
  if (!desc) { set_error("desc is null"); return IU_B200_EINVAL; }
  return create_from_arrays(desc->cell_type, desc->n_points, desc->n_cells, desc->points, desc->cells, desc->neighbors,
                            desc->n_point_data, desc->point_data, desc->n_cell_data, desc->cell_data,
                            desc->n_icell_data, desc->icell_data, desc->device, grid);
}

int iu_b200_grid_from_binda(const char* filename, double coord_scale_factor, int device, iu_b200_grid** grid) {
  if (!filename || !grid) { set_error("null argument"); return IU_B200_EINVAL; }
  HostMesh m;
  try {
    if (!read_binda_mesh(filename, coord_scale_factor, m)) return IU_B200_EIO;
  } catch (const std::bad_alloc&) {
    set_error("out of host memory");
    return IU_B200_ENOMEM;
  }
  int rc = create_from_arrays(m.cell_type, m.n_points, m.n_cells, m.points.data(), m.cells.data(), m.neighbors.data(),
                              (int)m.point_names.size(), m.point_data.data(), (int)m.cell_names.size(),
                              m.cell_data.data(), (int)m.icell_names.size(), m.icell_data.data(), device, grid);
  if (rc != IU_B200_OK) return rc;
  (*grid)->point_names = m.point_names;
  (*grid)->cell_names = m.cell_names;
  (*grid)->icell_names = m.icell_names;
  return IU_B200_OK;
}

int iu_b200_grid_replicate(iu_b200_grid* g, int n_devices, const int* devices) {
  if (!g) { set_error("null grid"); return IU_B200_EINVAL; }
  if (!check_device_available()) return IU_B200_ECUDA;
  int visible = 0;
  cudaGetDeviceCount(&visible);
  std::vector<int> want;
  if (n_devices <= 0 || !devices) { for (int d = 0; d < visible; ++d) want.push_back(d); }   // every visible GPU
  else want.assign(devices, devices + n_devices);
  std::lock_guard<std::mutex> lock(g->mu);
  return guarded([&]() {
    const auto t0 = std::chrono::steady_clock::now();
    for (int dev : want) {
      if (dev < 0 || dev >= visible) { set_error("device index outside 0..device count - 1"); return (int)IU_B200_EINVAL; }
      bool have = dev == g->device;
      for (iu_b200_grid* r : g->replicas) have = have || r->device == dev;
      // test hook: a second replica on a GPU that already has one exercises the clone and the slab threads on a one-GPU box
      const bool allow_same = std::getenv("IU_B200_ALLOW_SAME_DEVICE_REPLICAS") != nullptr;
      if (have && !(allow_same && n_devices > 0 && devices)) continue;
      { DeviceGuard guard(g->device); CU(cudaDeviceSynchronize()); }
      g->replicas.push_back(clone_to_device(g, dev));
    }
    g->replicate_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return (int)IU_B200_OK;
  });
}

int iu_b200_grid_create_multi(const iu_b200_grid_desc* desc, int n_devices, const int* devices, iu_b200_grid** grid) {
  if (!desc || !grid) { set_error("null argument"); return IU_B200_EINVAL; }
  iu_b200_grid_desc d0 = *desc;
  if (n_devices > 0 && devices) d0.device = devices[0];
  int rc = iu_b200_grid_create(&d0, grid);
  if (rc != IU_B200_OK) return rc;
  rc = iu_b200_grid_replicate(*grid, n_devices, devices);
  if (rc != IU_B200_OK) { iu_b200_grid_destroy(*grid); *grid = nullptr; }
  return rc;
}

int iu_b200_grid_from_binda_multi(const char* filename, double coord_scale_factor, int n_devices, const int* devices,
                                  iu_b200_grid** grid) {
  int rc = iu_b200_grid_from_binda(filename, coord_scale_factor, n_devices > 0 && devices ? devices[0] : 0, grid);
  if (rc != IU_B200_OK) return rc;
  rc = iu_b200_grid_replicate(*grid, n_devices, devices);
  if (rc != IU_B200_OK) { iu_b200_grid_destroy(*grid); *grid = nullptr; }
  return rc;
}

int iu_b200_device_count(const iu_b200_grid* g) { return g ? 1 + (int)g->replicas.size() : 0; }

iu_b200_grid* iu_b200_grid_replica(iu_b200_grid* g, int i) {
  if (!g || i < 0 || i > (int)g->replicas.size()) return nullptr;
  return i == 0 ? g : g->replicas[i - 1];
}

int iu_b200_grid_destroy(iu_b200_grid* g) {
  if (!g) return IU_B200_OK;
  for (iu_b200_grid* r : g->replicas) iu_b200_grid_destroy(r);
  g->replicas.clear();
  {
    DeviceGuard guard(g->device);
    cudaDeviceSynchronize();
    free_device_grid(g->dg);
    for (Slot& sl : g->slot) {
      if (sl.in_done) cudaEventDestroy(sl.in_done);
      if (sl.comp_done) cudaEventDestroy(sl.comp_done);
      if (sl.out_done) cudaEventDestroy(sl.out_done);
      sl.r.release(); sl.r1.release(); sl.out_f64.release();
      sl.i_a.release(); sl.i_b.release(); sl.i_c.release(); sl.i_d.release();
    }
    g->keys.release(); g->keys_alt.release(); g->idx.release(); g->idx_alt.release(); g->sort_temp.release();
    g->tr_seed.release(); g->tr_field.release(); g->tr_cell.release(); g->walk_sum.release();
    if (g->h_walk_sum) cudaFreeHost(g->h_walk_sum);
    if (g->walk_ev) cudaEventDestroy(g->walk_ev);
    if (g->ws_done) cudaEventDestroy(g->ws_done);
    g->idx_alt2.release();
    if (g->s_sort) cudaStreamDestroy(g->s_sort);
    if (g->ev_enter) cudaEventDestroy(g->ev_enter);
    for (int k = 0; k < 2; ++k) {
      if (g->ev_sorted[k]) cudaEventDestroy(g->ev_sorted[k]);
      if (g->ev_queried[k]) cudaEventDestroy(g->ev_queried[k]);
    }
    if (g->s_in) cudaStreamDestroy(g->s_in);
    if (g->s_comp) cudaStreamDestroy(g->s_comp);
    if (g->s_out) cudaStreamDestroy(g->s_out);
    if (g->h_counters) cudaFreeHost(g->h_counters);
    if (g->h_stage) cudaFreeHost(g->h_stage);
    collect_timings(g);
    for (cudaEvent_t e : g->event_pool) cudaEventDestroy(e);
  }
  delete g;
  return IU_B200_OK;
}

int iu_b200_grid_info(const iu_b200_grid* g, int32_t info[8]) {
  if (!g || !info) { set_error("null argument"); return IU_B200_EINVAL; }
  info[0] = g->dg.cell_type; info[1] = g->dg.n_points; info[2] = g->dg.n_cells; info[3] = g->dg.ppc;
  info[4] = g->dg.n_pd; info[5] = g->dg.n_cd; info[6] = g->dg.n_icd; info[7] = g->device;
  return IU_B200_OK;
}

int iu_b200_grid_bounds(const iu_b200_grid* g, double rmin[3], double rmax[3]) {
  if (!g || !rmin || !rmax) { set_error("null argument"); return IU_B200_EINVAL; }
  for (int k = 0; k < 3; ++k) { rmin[k] = g->dg.rmin[k]; rmax[k] = g->dg.rmax[k]; }
  return IU_B200_OK;
}

static int find_name(const std::vector<std::string>& names, const char* name) {
  if (!name) return -1;
  for (size_t i = 0; i < names.size(); ++i)
    if (names[i] == name) return (int)i + 1;
  return -1;
}
int iu_b200_get_point_data_index(const iu_b200_grid* g, const char* name) { return g ? find_name(g->point_names, name) : -1; }
int iu_b200_get_cell_data_index(const iu_b200_grid* g, const char* name) { return g ? find_name(g->cell_names, name) : -1; }
int iu_b200_get_icell_data_index(const iu_b200_grid* g, const char* name) { return g ? find_name(g->icell_names, name) : -1; }
int iu_b200_set_data_name(iu_b200_grid* g, int kind, int i_var, const char* name) {
  if (!g || !name || kind < 0 || kind > 2) { set_error("bad argument"); return IU_B200_EINVAL; }
  std::lock_guard<std::mutex> lock(g->mu);
  std::vector<std::string>& names = kind == 0 ? g->point_names : kind == 1 ? g->cell_names : g->icell_names;
  if (i_var < 1 || i_var > (int)names.size()) { set_error("column outside 1..n"); return IU_B200_EINVAL; }
  names[i_var - 1] = name;
  return on_replicas(g, IU_B200_OK, [&](iu_b200_grid* r) { return iu_b200_set_data_name(r, kind, i_var, name); });
}
const char* iu_b200_point_data_name(const iu_b200_grid* g, int i_var) {
  if (!g || i_var < 1 || i_var > (int)g->point_names.size()) return nullptr;
  return g->point_names[i_var - 1].c_str();
}
const char* iu_b200_cell_data_name(const iu_b200_grid* g, int i_var) {
  if (!g || i_var < 1 || i_var > (int)g->cell_names.size()) return nullptr;
  return g->cell_names[i_var - 1].c_str();
}
const char* iu_b200_icell_data_name(const iu_b200_grid* g, int i_var) {
  if (!g || i_var < 1 || i_var > (int)g->icell_names.size()) return nullptr;
  return g->icell_names[i_var - 1].c_str();
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// data columns, options, statistics
// ------------------------------------------------------------------------------------------------
namespace {

// grow a column-major (n_rows, n) device table by one zero column
template <class T>
void append_zero_column(T*& tab, int& n, int& cap, size_t n_rows) {
  if (n + 1 > cap) {
    const int new_cap = std::max(4, cap * 2);
    T* q = nullptr;
    CU(cudaMalloc((void**)&q, n_rows * new_cap * sizeof(T)));
    if (n > 0) CU(cudaMemcpy(q, tab, n_rows * n * sizeof(T), cudaMemcpyDeviceToDevice));
    cudaFree(tab);
    tab = q;
    cap = new_cap;
  }
  CU(cudaMemset(tab + n_rows * n, 0, n_rows * sizeof(T)));
  ++n;
}

}  // namespace

extern "C" {

int iu_b200_add_point_data(iu_b200_grid* g, const char* name, int* i_var) {
  if (!g || !name) { set_error("null argument"); return IU_B200_EINVAL; }
  std::lock_guard<std::mutex> lock(g->mu);
  int rc = guarded([&]() {
    DeviceGuard guard(g->device);
    DeviceGrid& d = g->dg;
    CU(cudaDeviceSynchronize());
    if (d.n_pd + 1 > d.pd_stride || !d.pd) {
      const int ns = round_up_stride(d.n_pd + 1);
      double* q = nullptr;
      CU(cudaMalloc((void**)&q, (size_t)d.n_points * ns * 8));
      if (d.pd) g->n_launches += launch_repack_point_data(d.pd, d.pd_stride, q, ns, d.n_pd, d.n_points, nullptr);
      else CU(cudaMemset(q, 0, (size_t)d.n_points * ns * 8));
      CU(cudaDeviceSynchronize());
      cudaFree(d.pd);
      d.pd = q;
      d.pd_stride = ns;
    } else {
      // the padding column may hold stale values of a column written before: zero it
      std::vector<double> z((size_t)d.n_points, 0.0);
      DevBuf<double> col;
      col.alloc(z.size());
      CU(cudaMemcpy(col.p, z.data(), z.size() * 8, cudaMemcpyHostToDevice));
      g->n_launches += launch_pack_point_data(d, col.p, d.n_pd, 1, nullptr);
      CU(cudaDeviceSynchronize());
    }
    ++d.n_pd;
    g->point_names.push_back(name);
    if (i_var) *i_var = d.n_pd;
    return IU_B200_OK;
  });
  return on_replicas(g, rc, [&](iu_b200_grid* r) { return iu_b200_add_point_data(r, name, nullptr); });
}

int iu_b200_add_cell_data(iu_b200_grid* g, const char* name, int* i_var) {
  if (!g || !name) { set_error("null argument"); return IU_B200_EINVAL; }
  std::lock_guard<std::mutex> lock(g->mu);
  int rc = guarded([&]() {
    DeviceGuard guard(g->device);
    CU(cudaDeviceSynchronize());
    append_zero_column(g->dg.cd, g->dg.n_cd, g->dg.cd_cap, (size_t)g->dg.n_cells);
    g->cell_names.push_back(name);
    if (i_var) *i_var = g->dg.n_cd;
    return IU_B200_OK;
  });
  return on_replicas(g, rc, [&](iu_b200_grid* r) { return iu_b200_add_cell_data(r, name, nullptr); });
}

int iu_b200_add_icell_data(iu_b200_grid* g, const char* name, int* i_var) {
  if (!g || !name) { set_error("null argument"); return IU_B200_EINVAL; }
  std::lock_guard<std::mutex> lock(g->mu);
  int rc = guarded([&]() {
    DeviceGuard guard(g->device);
    CU(cudaDeviceSynchronize());
    append_zero_column(g->dg.icd, g->dg.n_icd, g->dg.icd_cap, (size_t)g->dg.n_cells);
    g->icell_names.push_back(name);
    if (i_var) *i_var = g->dg.n_icd;
    return IU_B200_OK;
  });
  return on_replicas(g, rc, [&](iu_b200_grid* r) { return iu_b200_add_icell_data(r, name, nullptr); });
}

int iu_b200_update_point_data(iu_b200_grid* g, int col0, int n_cols, const double* data) {
  if (!g || !data) { set_error("null argument"); return IU_B200_EINVAL; }
  if (col0 < 1 || n_cols < 1 || col0 + n_cols - 1 > g->dg.n_pd) { set_error("point_data column range outside 1..n_point_data"); return IU_B200_EINVAL; }
  std::lock_guard<std::mutex> lock(g->mu);
  int rc = guarded([&]() {
    DeviceGuard guard(g->device);
    DevBuf<double> cols;
    cols.alloc((size_t)g->dg.n_points * n_cols);
    CU(cudaMemcpy(cols.p, data, cols.n * 8, cudaMemcpyHostToDevice));
    g->n_launches += launch_pack_point_data(g->dg, cols.p, col0 - 1, n_cols, nullptr);
    CU(cudaDeviceSynchronize());
    return IU_B200_OK;
  });
  return on_replicas(g, rc, [&](iu_b200_grid* r) { return iu_b200_update_point_data(r, col0, n_cols, data); });
}

int iu_b200_update_cell_data(iu_b200_grid* g, int col0, int n_cols, const double* data) {
  if (!g || !data) { set_error("null argument"); return IU_B200_EINVAL; }
  if (col0 < 1 || n_cols < 1 || col0 + n_cols - 1 > g->dg.n_cd) { set_error("cell_data column range outside 1..n_cell_data"); return IU_B200_EINVAL; }
  std::lock_guard<std::mutex> lock(g->mu);
  int rc = guarded([&]() {
    DeviceGuard guard(g->device);
    const size_t nc = (size_t)g->dg.n_cells;
    CU(cudaMemcpy(g->dg.cd + nc * (col0 - 1), data, nc * n_cols * 8, cudaMemcpyHostToDevice));
    return IU_B200_OK;
  });
  return on_replicas(g, rc, [&](iu_b200_grid* r) { return iu_b200_update_cell_data(r, col0, n_cols, data); });
}

int iu_b200_update_icell_data(iu_b200_grid* g, int col0, int n_cols, const int32_t* data) {
  if (!g || !data) { set_error("null argument"); return IU_B200_EINVAL; }
  if (col0 < 1 || n_cols < 1 || col0 + n_cols - 1 > g->dg.n_icd) { set_error("icell_data column range outside 1..n_icell_data"); return IU_B200_EINVAL; }
  std::lock_guard<std::mutex> lock(g->mu);
  int rc = guarded([&]() {
    DeviceGuard guard(g->device);
    const size_t nc = (size_t)g->dg.n_cells;
    CU(cudaMemcpy(g->dg.icd + nc * (col0 - 1), data, nc * n_cols * 4, cudaMemcpyHostToDevice));
    return IU_B200_OK;
  });
  return on_replicas(g, rc, [&](iu_b200_grid* r) { return iu_b200_update_icell_data(r, col0, n_cols, data); });
}

int iu_b200_download_point_data(iu_b200_grid* g, int col0, int n_cols, double* data) {
  if (!g || !data) { set_error("null argument"); return IU_B200_EINVAL; }
  if (col0 < 1 || n_cols < 1 || col0 + n_cols - 1 > g->dg.n_pd) { set_error("point_data column range outside 1..n_point_data"); return IU_B200_EINVAL; }
  std::lock_guard<std::mutex> lock(g->mu);
  return guarded([&]() {
    DeviceGuard guard(g->device);
    DevBuf<double> cols;
    cols.alloc((size_t)g->dg.n_points * n_cols);
    CU(cudaDeviceSynchronize());
    g->n_launches += launch_unpack_point_data(g->dg, cols.p, col0 - 1, n_cols, nullptr);
    CU(cudaMemcpy(data, cols.p, cols.n * 8, cudaMemcpyDeviceToHost));
    return IU_B200_OK;
  });
}

int iu_b200_download_cell_data(iu_b200_grid* g, int col0, int n_cols, double* data) {
  if (!g || !data) { set_error("null argument"); return IU_B200_EINVAL; }
  if (col0 < 1 || n_cols < 1 || col0 + n_cols - 1 > g->dg.n_cd) { set_error("cell_data column range outside 1..n_cell_data"); return IU_B200_EINVAL; }
  std::lock_guard<std::mutex> lock(g->mu);
  return guarded([&]() {
    DeviceGuard guard(g->device);
    const size_t nc = (size_t)g->dg.n_cells;
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(data, g->dg.cd + nc * (col0 - 1), nc * n_cols * 8, cudaMemcpyDeviceToHost));
    return IU_B200_OK;
  });
}

int iu_b200_download_icell_data(iu_b200_grid* g, int col0, int n_cols, int32_t* data) {
  if (!g || !data) { set_error("null argument"); return IU_B200_EINVAL; }
  if (col0 < 1 || n_cols < 1 || col0 + n_cols - 1 > g->dg.n_icd) { set_error("icell_data column range outside 1..n_icell_data"); return IU_B200_EINVAL; }
  std::lock_guard<std::mutex> lock(g->mu);
  return guarded([&]() {
    DeviceGuard guard(g->device);
    const size_t nc = (size_t)g->dg.n_cells;
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpy(data, g->dg.icd + nc * (col0 - 1), nc * n_cols * 4, cudaMemcpyDeviceToHost));
    return IU_B200_OK;
  });
}

static int set_option_one(iu_b200_grid* g, int option, int64_t value);
int iu_b200_set_option(iu_b200_grid* g, int option, int64_t value) {
  if (!g) { set_error("null grid"); return IU_B200_EINVAL; }
  return on_replicas(g, set_option_one(g, option, value), [&](iu_b200_grid* r) { return set_option_one(r, option, value); });
}
static int set_option_one(iu_b200_grid* g, int option, int64_t value) {
  switch (option) {
    case IU_B200_OPT_SORT_QUERIES:
      if (value < 0 || value > 3) break;
      g->sort_mode = (int)value;
      return IU_B200_OK;
    case IU_B200_OPT_MAX_WALK_STEPS:
      if (value < 1 || value > 0x7ffffffe) break;
      g->max_walk_steps = value;
      return IU_B200_OK;
    case IU_B200_OPT_HOST_CHUNK:
      if (value < 1 || value > (1LL << 30)) break;
      g->host_chunk = value;
      return IU_B200_OK;
    case IU_B200_OPT_DEV_BATCH:
      if (value < 1 || value > (1LL << 31)) break;
      g->dev_batch = value;
      return IU_B200_OK;
    case IU_B200_OPT_DYNAMIC_TILES:
      if (value < 0 || value > 2) break;
      g->dynamic_tiles = (int)value;
      return IU_B200_OK;
    case IU_B200_OPT_OVERLAP_SORT:
      if (value < 0 || value > 1) break;
      g->overlap_sort = (int)value;
      return IU_B200_OK;
    case IU_B200_OPT_TILES_PER_WARP:
      if (value < 1 || value > (1 << 20)) break;
      g->tiles_per_warp = (int)value;
      return IU_B200_OK;
    case IU_B200_OPT_TIME_KERNELS:
      if (value < 0 || value > 1) break;
      g->time_kernels = value != 0;
      return IU_B200_OK;
    case IU_B200_OPT_NN_VOXEL_DENSITY: {
      if (value < 0 || value > 6400) break;
      std::lock_guard<std::mutex> lock(g->mu);
      return guarded([&]() {
        DeviceGuard guard(g->device);
        g->voxel_density = (int)value;
        build_voxel_table(g);
        return IU_B200_OK;
      });
    }
    default:
      set_error("unknown option");
      return IU_B200_EINVAL;
  }
  set_error("option value out of range");
  return IU_B200_EINVAL;
}

int iu_b200_get_stats(iu_b200_grid* g, int64_t stats[IU_B200_N_STATS]) {
  if (!g || !stats) { set_error("null argument"); return IU_B200_EINVAL; }
  for (int i = 0; i < IU_B200_N_STATS; ++i) stats[i] = 0;
  std::lock_guard<std::mutex> lock(g->mu);
  return guarded([&]() {
    DeviceGuard guard(g->device);
    unsigned long long c[CNT_N];
    CU(cudaMemcpy(c, g->dg.counters, sizeof(c), cudaMemcpyDeviceToHost));
    stats[IU_B200_STAT_KERNEL_LAUNCHES] = g->n_launches;
    stats[IU_B200_STAT_WALK_LIMIT_HITS] = (int64_t)c[CNT_WALK_LIMIT];
    stats[IU_B200_STAT_QUERIES] = g->n_queries;
    stats[IU_B200_STAT_WORKSPACE_BYTES] = (int64_t)workspace_bytes(g);
    stats[IU_B200_STAT_MESH_BYTES] = (int64_t)(g->mesh_bytes + (size_t)g->dg.n_vox * VOX_K * 4);
    stats[IU_B200_STAT_NN_VOXELS] = g->dg.n_vox;
    stats[IU_B200_STAT_NN_VOXELS_TREE] = g->vox_flagged;
    stats[IU_B200_STAT_NN_VOXEL_ENTRIES] = g->vox_entries;
    stats[IU_B200_STAT_NN_TREE_QUERIES] = (int64_t)c[CNT_NN_TREE];
    stats[IU_B200_STAT_NN_VOXEL_BUILD_US] = (int64_t)(g->vox_build_ms * 1e3);
    stats[IU_B200_STAT_TREE_DEPTH] = g->tree_depth;
    stats[IU_B200_STAT_NOT_FOUND] = (int64_t)c[CNT_NOT_FOUND];
    collect_timings(g);
    stats[IU_B200_STAT_QUERY_KERNEL_NS] = (int64_t)g->timed_ns[0];
    stats[IU_B200_STAT_QUERY_KERNEL_LAUNCHES] = g->timed_launches[0];
    stats[IU_B200_STAT_SORT_NS] = (int64_t)g->timed_ns[1];
    stats[IU_B200_STAT_DEVICES] = 1 + (int64_t)g->replicas.size();
    stats[IU_B200_STAT_REPLICATE_US] = (int64_t)(g->replicate_ms * 1e3);
    stats[IU_B200_STAT_BOUNDS_ERRORS] = (int64_t)bounds_errors();
    stats[IU_B200_STAT_H2D_SKIPPED_BYTES] = (int64_t)g->h2d_skipped_bytes;
    stats[IU_B200_STAT_TREE_BUILD_US] = (int64_t)(g->tree_build_ms * 1e3);
    // a multi-GPU handle reports the sums over its devices for the counters that add up
    for (iu_b200_grid* r : g->replicas) {
      int64_t rs[IU_B200_N_STATS];
      const int rc = iu_b200_get_stats(r, rs);
      if (rc != IU_B200_OK) return rc;
      for (int k : {IU_B200_STAT_KERNEL_LAUNCHES, IU_B200_STAT_WALK_LIMIT_HITS, IU_B200_STAT_QUERIES, IU_B200_STAT_WORKSPACE_BYTES,
                    IU_B200_STAT_MESH_BYTES, IU_B200_STAT_NN_TREE_QUERIES, IU_B200_STAT_NOT_FOUND, IU_B200_STAT_QUERY_KERNEL_NS,
                    IU_B200_STAT_QUERY_KERNEL_LAUNCHES, IU_B200_STAT_SORT_NS, IU_B200_STAT_H2D_SKIPPED_BYTES})
        stats[k] += rs[k];
      if (rs[IU_B200_STAT_BOUNDS_ERRORS] > 0) stats[IU_B200_STAT_BOUNDS_ERRORS] += rs[IU_B200_STAT_BOUNDS_ERRORS];
    }
    return (int)IU_B200_OK;
  });
}

int iu_b200_reset_stats(iu_b200_grid* g) {
  if (!g) { set_error("null grid"); return IU_B200_EINVAL; }
  std::lock_guard<std::mutex> lock(g->mu);
  int rc = guarded([&]() {
    DeviceGuard guard(g->device);
    CU(cudaDeviceSynchronize());
    unsigned long long c[CNT_N];
    CU(cudaMemcpy(c, g->dg.counters, sizeof(c), cudaMemcpyDeviceToHost));
    c[CNT_WALK_LIMIT] = 0;
    c[CNT_NOT_FOUND] = 0;
    c[CNT_NN_TREE] = 0;
    c[CNT_EARLY] = 0;   // the hint-error counter is kept: iu_b200_check_dev compares it with `seen`
    CU(cudaMemcpy(g->dg.counters, c, sizeof(c), cudaMemcpyHostToDevice));
    for (int i = 0; i < CNT_N; ++i) g->seen[i] = c[i];
    g->n_launches = 0;
    g->n_queries = 0;
    g->h2d_skipped_bytes = 0;
    collect_timings(g);
    g->timed_ns[0] = g->timed_ns[1] = 0;
    g->timed_launches[0] = g->timed_launches[1] = 0;
    return IU_B200_OK;
  });
  return on_replicas(g, rc, [&](iu_b200_grid* r) { return iu_b200_reset_stats(r); });
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// the hot path on device buffers
// ------------------------------------------------------------------------------------------------
namespace {

bool want_sort(const iu_b200_grid* g, int mode, int64_t m) {
  if (mode == QM_INTERP_ICELL) return false;
  if (g->sort_mode == 0) return false;
  if (g->sort_mode == 1) return m > 1;
  return m >= (1 << 15) && g->dg.n_cells >= (1 << 12);
}

// Runs one logical batch as sub-batches of at most dev_batch queries; each sub-batch is optionally
// Morton-sorted (key, index) and processed through the permutation.
// Sub-batch view of a logical batch
QueryArgs sub_batch(const QueryArgs& a, int64_t off, int64_t m, size_t out_row) {
  QueryArgs b = a;
  b.n = m;
  b.r = a.r + 3 * off;
  b.i_cell = a.i_cell ? a.i_cell + off : nullptr;
  b.out_f64 = a.out_f64 ? a.out_f64 + out_row * off : nullptr;
  b.out_i32 = a.out_i32 ? a.out_i32 + off : nullptr;
  b.out2_f64 = a.out2_f64 ? a.out2_f64 + off : nullptr;
  b.perm = nullptr;
  return b;
}

// Several sorted sub-batches: the curve keys + radix sort of sub-batch k+1 run on a second, high-priority stream while
// k_query works on sub-batch k.  The sort is HBM-bound, k_query is L1/latency-bound, so they share the SMs well - but
// only if k_query leaves room: its warps retire after tiles_per_warp tiles (short-lived CTAs instead of a persistent
// grid), and the block scheduler fills freed slots with the sort's CTAs first.  Two permutation buffers alternate.
void run_query_overlapped(iu_b200_grid* g, const QueryArgs& a, cudaStream_t s, size_t out_row) {
  const int64_t n = a.n, step = g->dev_batch;
  const int end_bit = morton_end_bit(g->dg);
  g->keys.ensure(step); g->keys_alt.ensure(step); g->idx.ensure(step); g->idx_alt.ensure(step); g->idx_alt2.ensure(step);
  g->sort_temp.ensure(sort_temp_bytes(step, end_bit));
  if (!g->s_sort) {
    int lo = 0, hi = 0;
    CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CU(cudaStreamCreateWithPriority(&g->s_sort, cudaStreamNonBlocking, hi));
    CU(cudaEventCreateWithFlags(&g->ev_enter, cudaEventDisableTiming));
    for (int k = 0; k < 2; ++k) {
      CU(cudaEventCreateWithFlags(&g->ev_sorted[k], cudaEventDisableTiming));
      CU(cudaEventCreateWithFlags(&g->ev_queried[k], cudaEventDisableTiming));
    }
  }
  uint32_t* perm[2] = {g->idx_alt.p, g->idx_alt2.p};
  CU(cudaEventRecord(g->ev_enter, s));                       // the caller's inputs are ready from here on
  CU(cudaStreamWaitEvent(g->s_sort, g->ev_enter, 0));
  const int64_t n_sub = (n + step - 1) / step;
  auto enqueue_sort = [&](int64_t k) {
    const int64_t off = k * step, m = std::min(step, n - off);
    if (k >= 2) CU(cudaStreamWaitEvent(g->s_sort, g->ev_queried[k & 1], 0));   // the query that read this buffer last
    TimedScope ts(g, g->s_sort, 1);
    g->n_launches += launch_morton_sort(g->dg, m, a.r + 3 * off, a.mode == QM_NEAREST || !a.i_cell ? nullptr : a.i_cell + off, g->keys.p,
                                        g->keys_alt.p, g->idx.p, perm[k & 1], g->sort_temp.p, g->sort_temp.n, nullptr, g->s_sort);
    ts.stop();
    CU(cudaEventRecord(g->ev_sorted[k & 1], g->s_sort));
  };
  enqueue_sort(0);
  for (int64_t k = 0; k < n_sub; ++k) {
    const int64_t off = k * step, m = std::min(step, n - off);
    if (k + 1 < n_sub) enqueue_sort(k + 1);
    QueryArgs b = sub_batch(a, off, m, out_row);
    b.perm = perm[k & 1];
    b.dynamic = g->dynamic_tiles;
    b.tiles_per_warp = k + 1 < n_sub ? g->tiles_per_warp : 0;   // nothing runs beside the last one
    CU(cudaStreamWaitEvent(s, g->ev_sorted[k & 1], 0));
    TimedScope tq(g, s, 0);
    g->n_launches += launch_query(g->dg, b, s);
    tq.stop();
    CU(cudaEventRecord(g->ev_queried[k & 1], s));
    CU(cudaGetLastError());
  }
  g->n_queries += n;
}

void run_query(iu_b200_grid* g, QueryArgs a, cudaStream_t s) {
  WorkspaceScope ws(g, s);
  const int64_t n = a.n;
  const int64_t step = g->dev_batch;
  const size_t out_row = a.mode == QM_INTERP || a.mode == QM_INTERP_ICELL || a.mode == QM_INTERP_CELL ? (size_t)a.n_vars : 1;
  if (g->overlap_sort && g->dynamic_tiles == 1 && g->sort_mode != 3 && n > step && want_sort(g, a.mode, std::min(step, n - (n - 1) / step * step))) {
    run_query_overlapped(g, a, s, out_row);
    return;
  }
  for (int64_t off = 0; off < n; off += step) {
    const int64_t m = std::min(step, n - off);
    QueryArgs b = a;
    b.n = m;
    b.r = a.r + 3 * off;
    b.i_cell = a.i_cell ? a.i_cell + off : nullptr;
    b.out_f64 = a.out_f64 ? a.out_f64 + out_row * off : nullptr;
    b.out_i32 = a.out_i32 ? a.out_i32 + off : nullptr;
    b.out2_f64 = a.out2_f64 ? a.out2_f64 + off : nullptr;
    b.perm = nullptr;
    if (g->sort_mode == 3) {   // the caller's queries are already spatially ordered: ticket tiles without a permutation
      b.dynamic = g->dynamic_tiles;
    } else if (want_sort(g, a.mode, m)) {
      const int end_bit = morton_end_bit(g->dg);
      g->keys.ensure(m); g->keys_alt.ensure(m); g->idx.ensure(m); g->idx_alt.ensure(m);
      g->sort_temp.ensure(sort_temp_bytes(m, end_bit));
      TimedScope ts(g, s, 1);
      g->n_launches += launch_morton_sort(g->dg, m, b.r, a.mode == QM_NEAREST ? nullptr : b.i_cell, g->keys.p, g->keys_alt.p,
                                          g->idx.p, g->idx_alt.p, g->sort_temp.p, g->sort_temp.n, nullptr, s);
      ts.stop();
      b.perm = g->idx_alt.p;
      b.dynamic = g->dynamic_tiles;
    }
    TimedScope tq(g, s, 0);
    g->n_launches += launch_query(g->dg, b, s);
    tq.stop();
    CU(cudaGetLastError());
  }
  g->n_queries += n;
}

// iu_get_cell_through_neighbors for a batch: sub-batches of at most dev_batch rays.  Only with
// IU_B200_OPT_SORT_QUERIES = 1 are they processed in Morton order of the start positions: measured on
// C4 (100M rays), the 13 scattered accesses per ray through the permutation cost more than the
// locality gains below ~2 cells per ray (18.5 -> 29.7 ms at 0.1 h) and pay off above (71 -> 32.6 ms at 1 h).
// huge(1) (what the reference's own callers pass, :431, :1122) would overflow the do-loop counter on a walk that never
// terminates (n_steps = max_steps + 1 after the loop): the cap keeps that sum representable
constexpr int MAX_STEPS_CAP = 0x7ffffffe;
constexpr int64_t WALK_AUTO_MIN_RAYS = 1 << 20;
constexpr int WALK_AUTO_SAMPLES = 4096;
constexpr double WALK_AUTO_CELLS = 0.55;

void run_walk(iu_b200_grid* g, WalkArgs a, cudaStream_t s) {
  WorkspaceScope ws(g, s);
  const int64_t n = a.n;
  const int64_t step = g->dev_batch;
  // auto: sorting pays once the rays cross more than about one cell (profiles/README.md, C4 table: unsorted
  // 0.13 + 0.31 d ns per ray, sorted 0.29 + 0.02 d, d = ray length in mean cell sizes).  d comes from a sample of a
  // call's rays, read back asynchronously: the first large call waits for its own sample, later calls use the latest
  // sample that has arrived (tracers call with similar rays step after step) and never synchronise the stream.
  bool auto_sort = false;
  if (g->sort_mode == 2 && n >= WALK_AUTO_MIN_RAYS) {
    const DeviceGrid& dg = g->dg;
    double ext = 1.0; int nd = 0;
    for (int k = 0; k < 3; ++k) if (dg.rmax[k] > dg.rmin[k]) { ext *= dg.rmax[k] - dg.rmin[k]; ++nd; }
    if (nd >= 2 && dg.n_cells > 0) {
      const double h_cell = std::pow(ext / dg.n_cells, 1.0 / nd);
      if (!g->h_walk_sum) {
        CU(cudaHostAlloc((void**)&g->h_walk_sum, sizeof(double), cudaHostAllocDefault));
        CU(cudaEventCreateWithFlags(&g->walk_ev, cudaEventDisableTiming));
        g->walk_sum.ensure(1);
      }
      auto take = [&]() {
        g->walk_sort = *g->h_walk_sum / WALK_AUTO_SAMPLES > WALK_AUTO_CELLS * h_cell;
        g->walk_ev_pending = false;
        g->walk_measured = true;
      };
      if (g->walk_ev_pending && cudaEventQuery(g->walk_ev) == cudaSuccess) take();
      if (!g->walk_ev_pending) {
        g->n_launches += launch_ray_length_sample(n, WALK_AUTO_SAMPLES, a.r0, a.r1, g->walk_sum.p, s);
        CU(cudaMemcpyAsync(g->h_walk_sum, g->walk_sum.p, sizeof(double), cudaMemcpyDeviceToHost, s));
        CU(cudaEventRecord(g->walk_ev, s));
        g->walk_ev_pending = true;
      }
      if (!g->walk_measured) { CU(cudaEventSynchronize(g->walk_ev)); take(); }
      auto_sort = g->walk_sort;
    }
  }
  for (int64_t off = 0; off < n; off += step) {
    const int64_t m = std::min(step, n - off);
    WalkArgs b = a;
    b.n = m;
    b.r0 = a.r0 + 3 * off; b.r1 = a.r1 + 3 * off; b.ic0 = a.ic0 + off; b.ic1 = a.ic1 + off;
    b.r_p = a.r_p + 3 * off; b.n_steps = a.n_steps + off; b.status = a.status + off;
    b.perm = nullptr;
    const bool sort = m > 1 && (g->sort_mode == 1 || auto_sort);
    if (sort) {
      const int end_bit = morton_end_bit(g->dg);
      g->keys.ensure(m); g->keys_alt.ensure(m); g->idx.ensure(m); g->idx_alt.ensure(m);
      g->sort_temp.ensure(sort_temp_bytes(m, end_bit));
      TimedScope ts(g, s, 1);
      g->n_launches += launch_morton_sort(g->dg, m, b.r0, nullptr, g->keys.p, g->keys_alt.p, g->idx.p, g->idx_alt.p,
                                          g->sort_temp.p, g->sort_temp.n, nullptr, s);
      ts.stop();
      b.perm = g->idx_alt.p;
    }
    g->n_launches += launch_walk(g->dg, b, s);
    CU(cudaGetLastError());
  }
  g->n_queries += n;
}

int check_query_args(const iu_b200_grid* g, int64_t n, const void* r, int mode, int n_vars, const int32_t* i_vars, int i_var) {
  if (!g) { set_error("null grid"); return IU_B200_EINVAL; }
  if (n < 0) { set_error("negative query count"); return IU_B200_EINVAL; }
  if (n > 0 && !r) { set_error("null query array"); return IU_B200_EINVAL; }
  if (mode == QM_INTERP || mode == QM_INTERP_ICELL || mode == QM_INTERP_CELL) {
    if (n_vars < 1 || n_vars > MAX_VARS || !i_vars) { set_error("n_vars must be in 1..32"); return IU_B200_EINVAL; }
    for (int v = 0; v < n_vars; ++v)
      if (i_vars[v] < 1 || i_vars[v] > g->dg.n_pd) { set_error("i_vars: column outside 1..n_point_data"); return IU_B200_EINVAL; }
    if (mode == QM_INTERP_CELL && (i_var < 1 || i_var > g->dg.n_cd)) { set_error("i_var outside 1..n_cell_data"); return IU_B200_EINVAL; }
  } else if (mode == QM_CELL_F64) {
    if (i_var < 1 || i_var > g->dg.n_cd) { set_error("i_var outside 1..n_cell_data"); return IU_B200_EINVAL; }
  } else if (mode == QM_CELL_I32) {
    if (i_var < 1 || i_var > g->dg.n_icd) { set_error("i_var outside 1..n_icell_data"); return IU_B200_EINVAL; }
  }
  return IU_B200_OK;
}

QueryArgs make_args(const iu_b200_grid* g, int mode, int64_t n, const double* r, int n_vars, const int32_t* i_vars, int i_var,
                    int32_t* i_cell, double* out_f64, int32_t* out_i32, double* out2_f64 = nullptr) {
  QueryArgs a;
  a.mode = mode; a.n = n; a.r = r; a.i_cell = i_cell; a.out_f64 = out_f64; a.out_i32 = out_i32; a.out2_f64 = out2_f64;
  a.max_steps = (int)g->max_walk_steps;
  if (mode == QM_INTERP || mode == QM_INTERP_ICELL || mode == QM_INTERP_CELL) {
    a.n_vars = n_vars;
    for (int v = 0; v < n_vars; ++v) a.ivars[v] = i_vars[v] - 1;
    if (mode == QM_INTERP_CELL) a.cell_var = i_var - 1;
  } else if (mode == QM_CELL_F64 || mode == QM_CELL_I32) {
    a.n_vars = 1;
    a.ivars[0] = i_var - 1;
  }
  return a;
}

int query_dev(iu_b200_grid* g, int mode, int64_t n, const double* r, int n_vars, const int32_t* i_vars, int i_var,
              int32_t* i_cell, double* out_f64, int32_t* out_i32, void* stream, double* out2_f64 = nullptr) {
  int rc = check_query_args(g, n, r, mode, n_vars, i_vars, i_var);
  if (rc != IU_B200_OK || n == 0) return rc;
  std::lock_guard<std::mutex> lock(g->mu);
  return guarded([&]() {
    DeviceGuard guard(g->device);
    run_query(g, make_args(g, mode, n, r, n_vars, i_vars, i_var, i_cell, out_f64, out_i32, out2_f64), (cudaStream_t)stream);
    return IU_B200_OK;
  });
}

int walk_args_ok(const iu_b200_grid* g, int64_t n, int max_steps, const void* r0, const void* r1, const void* ic0,
                 const void* ic1, const void* r_p, const void* n_steps, const void* status, int mask) {
  if (!g) { set_error("null grid"); return IU_B200_EINVAL; }
  if (n < 0 || max_steps < 0) { set_error("negative count"); return IU_B200_EINVAL; }
  if (n > 0 && (!r0 || !r1 || !ic0 || !ic1 || !r_p || !n_steps || !status)) { set_error("null array"); return IU_B200_EINVAL; }
  if (mask < 0 || mask > g->dg.n_icd) { set_error("i_icell_mask outside 0..n_icell_data"); return IU_B200_EINVAL; }
  return IU_B200_OK;
}

}  // namespace

extern "C" {

int iu_b200_find_nearby_cell_batch_dev(iu_b200_grid* g, int64_t n, const double* r, int32_t* idx, void* stream) {
  return query_dev(g, QM_NEAREST, n, r, 0, nullptr, 0, nullptr, nullptr, idx, stream);
}
int iu_b200_get_cell_batch_dev(iu_b200_grid* g, int64_t n, const double* r, int32_t* i_cell, void* stream) {
  return query_dev(g, QM_CELL, n, r, 0, nullptr, 0, i_cell, nullptr, nullptr, stream);
}
int iu_b200_interpolate_at_batch_dev(iu_b200_grid* g, int64_t n, const double* r, int n_vars, const int32_t* i_vars,
                                     double* var_at_r, int32_t* i_cell, void* stream) {
  return query_dev(g, QM_INTERP, n, r, n_vars, i_vars, 0, i_cell, var_at_r, nullptr, stream);
}
int iu_b200_interpolate_at_icell_batch_dev(iu_b200_grid* g, int64_t n, const double* r, int n_vars, const int32_t* i_vars,
                                           double* var_at_r, const int32_t* i_cell, void* stream) {
  return query_dev(g, QM_INTERP_ICELL, n, r, n_vars, i_vars, 0, const_cast<int32_t*>(i_cell), var_at_r, nullptr, stream);
}
int iu_b200_interpolate_and_cell_scalar_at_batch_dev(iu_b200_grid* g, int64_t n, const double* r, int n_vars,
                                                     const int32_t* i_vars, int i_cell_var, double* var_at_r, double* cell_res,
                                                     int32_t* i_cell, void* stream) {
  if (n > 0 && (!var_at_r || !cell_res || !i_cell)) { set_error("null output array"); return IU_B200_EINVAL; }
  return query_dev(g, QM_INTERP_CELL, n, r, n_vars, i_vars, i_cell_var, i_cell, var_at_r, nullptr, stream, cell_res);
}
int iu_b200_get_cell_scalar_at_batch_dev(iu_b200_grid* g, int64_t n, const double* r, int i_var, int32_t* i_guess,
                                         double* res, void* stream) {
  return query_dev(g, QM_CELL_F64, n, r, 0, nullptr, i_var, i_guess, res, nullptr, stream);
}
int iu_b200_get_icell_scalar_at_batch_dev(iu_b200_grid* g, int64_t n, const double* r, int i_var, int32_t* i_guess,
                                          int32_t* res, void* stream) {
  return query_dev(g, QM_CELL_I32, n, r, 0, nullptr, i_var, i_guess, nullptr, res, stream);
}

int iu_b200_get_cell_through_neighbors_batch_dev(iu_b200_grid* g, int64_t n, int max_steps, const double* r0,
                                                 const double* r1, const int32_t* ic0, int32_t* ic1, double* r_p,
                                                 int32_t* n_steps, int32_t* status, int i_icell_mask, void* stream) {
  int rc = walk_args_ok(g, n, max_steps, r0, r1, ic0, ic1, r_p, n_steps, status, i_icell_mask);
  if (rc != IU_B200_OK || n == 0) return rc;
  std::lock_guard<std::mutex> lock(g->mu);
  return guarded([&]() {
    DeviceGuard guard(g->device);
    WalkArgs a;
    a.n = n; a.max_steps = std::min(max_steps, MAX_STEPS_CAP); a.r0 = r0; a.r1 = r1; a.ic0 = ic0; a.ic1 = ic1; a.r_p = r_p;
    a.n_steps = n_steps; a.status = status; a.mask_col = i_icell_mask - 1;
    run_walk(g, a, (cudaStream_t)stream);
    return IU_B200_OK;
  });
}

int iu_b200_interpolate_at_batch_dev_noalloc(iu_b200_grid* g, int64_t n, const double* r, int n_vars, const int32_t* i_vars,
                                             double* var_at_r, int32_t* i_cell, void* stream) {
  int rc = check_query_args(g, n, r, QM_INTERP, n_vars, i_vars, 0);
  if (rc != IU_B200_OK || n == 0) return rc;
  if (!var_at_r || !i_cell) { set_error("null output array"); return IU_B200_EINVAL; }
  if (n > NOALLOC_MAX_QUERIES) {
    set_error("iu_b200_interpolate_at_batch_dev_noalloc: n exceeds IU_B200_NOALLOC_MAX_QUERIES (1048576); split the batch or use "
              "iu_b200_interpolate_at_batch_dev");
    return IU_B200_EINVAL;
  }
  std::lock_guard<std::mutex> lock(g->mu);
  DeviceGuard guard(g->device);
  const int launches = launch_query_noalloc(g->dg, make_args(g, QM_INTERP, n, r, n_vars, i_vars, 0, i_cell, var_at_r, nullptr), (cudaStream_t)stream);
  if (launches < 0) return IU_B200_ECUDA;
  g->n_launches += launches;
  g->n_queries += n;
  return IU_B200_OK;
}

int iu_b200_check_dev(iu_b200_grid* g, void* stream) {
  if (!g) { set_error("null grid"); return IU_B200_EINVAL; }
  std::lock_guard<std::mutex> lock(g->mu);
  return guarded([&]() {
    DeviceGuard guard(g->device);
    CU(cudaStreamSynchronize((cudaStream_t)stream));
    unsigned long long c[CNT_N];
    CU(cudaMemcpy(c, g->dg.counters, sizeof(c), cudaMemcpyDeviceToHost));
    const bool bad = c[CNT_HINT_ERROR] != g->seen[CNT_HINT_ERROR];
    g->seen[CNT_HINT_ERROR] = c[CNT_HINT_ERROR];
    if (bad) {
      set_error("i_cell > ug%n_cells (or a start cell outside 1..n_cells)");
      return IU_B200_EHINT;
    }
    return IU_B200_OK;
  });
}

int iu_b200_generate_queries_dev(iu_b200_grid* g, int64_t n, int64_t first, uint64_t seed, const double lo[3],
                                 const double hi[3], double* r, void* stream) {
  if (!g || !lo || !hi || (n > 0 && !r) || n < 0) { set_error("bad argument"); return IU_B200_EINVAL; }
  return guarded([&]() {
    DeviceGuard guard(g->device);
    g->n_launches += launch_generate_queries(n, first, seed, lo, hi, r, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return IU_B200_OK;
  });
}

int iu_b200_export_geometry(iu_b200_grid* g, double* cell_points, double* cell_face_normals, double* cell_volume,
                            double* cell_centers) {
  if (!g) { set_error("null grid"); return IU_B200_EINVAL; }
  std::lock_guard<std::mutex> lock(g->mu);
  return guarded([&]() {
    DeviceGuard guard(g->device);
    const size_t nc = (size_t)g->dg.n_cells, ppc = (size_t)g->dg.ppc;
    DevBuf<double> cp, fn, vol, ce;
    if (cell_points) cp.alloc(nc * ppc * 3);
    if (cell_face_normals) fn.alloc(nc * ppc * 3);
    if (cell_volume) vol.alloc(nc);
    if (cell_centers) ce.alloc(nc * 3);
    g->n_launches += launch_export_geometry(g->dg, cp.p, fn.p, vol.p, ce.p, nullptr);
    CU(cudaDeviceSynchronize());
    if (cell_points) CU(cudaMemcpy(cell_points, cp.p, cp.n * 8, cudaMemcpyDeviceToHost));
    if (cell_face_normals) CU(cudaMemcpy(cell_face_normals, fn.p, fn.n * 8, cudaMemcpyDeviceToHost));
    if (cell_volume) CU(cudaMemcpy(cell_volume, vol.p, vol.n * 8, cudaMemcpyDeviceToHost));
    if (cell_centers) CU(cudaMemcpy(cell_centers, ce.p, ce.n * 8, cudaMemcpyDeviceToHost));
    return IU_B200_OK;
  });
}

int iu_b200_export_mesh(iu_b200_grid* g, double* points, int32_t* cells) {
  if (!g) { set_error("null grid"); return IU_B200_EINVAL; }
  std::lock_guard<std::mutex> lock(g->mu);
  return guarded([&]() {
    DeviceGuard guard(g->device);
    DevBuf<double> p;
    DevBuf<int32_t> c;
    if (cells) c.alloc((size_t)g->dg.ppc * g->dg.n_cells);
    CU(cudaDeviceSynchronize());
    if (cells) g->n_launches += launch_export_mesh(g->dg, nullptr, c.p, nullptr);
    CU(cudaDeviceSynchronize());
    if (points) CU(cudaMemcpy(points, g->dg.points, (size_t)3 * g->dg.n_points * 8, cudaMemcpyDeviceToHost));
    if (cells) CU(cudaMemcpy(cells, c.p, c.n * 4, cudaMemcpyDeviceToHost));
    return IU_B200_OK;
  });
}

int iu_b200_write_vtk(iu_b200_grid* g, const char* filename) {
  if (!g || !filename) { set_error("null argument"); return IU_B200_EINVAL; }
  const DeviceGrid& d = g->dg;
  std::vector<double> points((size_t)3 * d.n_points), pd((size_t)d.n_points * d.n_pd), cd((size_t)d.n_cells * d.n_cd);
  std::vector<int32_t> cells((size_t)d.ppc * d.n_cells), icd((size_t)d.n_cells * d.n_icd);
  int rc = iu_b200_export_mesh(g, points.data(), cells.data());
  if (rc == IU_B200_OK && d.n_pd > 0) rc = iu_b200_download_point_data(g, 1, d.n_pd, pd.data());
  if (rc == IU_B200_OK && d.n_cd > 0) rc = iu_b200_download_cell_data(g, 1, d.n_cd, cd.data());
  if (rc == IU_B200_OK && d.n_icd > 0) rc = iu_b200_download_icell_data(g, 1, d.n_icd, icd.data());
  if (rc != IU_B200_OK) return rc;
  auto named = [](const std::vector<std::string>& names, int n, const char* prefix, const void* data) {
    VtuColumns c;
    c.data = data;
    for (int v = 0; v < n; ++v)   // columns created without a name (iu_b200_grid_create) get "<prefix><column>"
      c.names.push_back(v < (int)names.size() && !names[v].empty() ? names[v] : prefix + std::to_string(v + 1));
    return c;
  };
  if (!write_vtu(filename, d.cell_type, d.n_points, d.n_cells, points.data(), cells.data(),
                 named(g->cell_names, d.n_cd, "cell_data_", cd.data()), named(g->icell_names, d.n_icd, "icell_data_", icd.data()),
                 named(g->point_names, d.n_pd, "point_data_", pd.data())))
    return IU_B200_EIO;
  return IU_B200_OK;
}

int iu_b200_write_vtk_arrays(const char* filename, int cell_type, int32_t n_points, int32_t n_cells, const double* points,
                             const int32_t* cells, int n_cell_data, const double* cell_data, const char* const* cell_names,
                             int n_icell_data, const int32_t* icell_data, const char* const* icell_names,
                             int n_point_data, const double* point_data, const char* const* point_names) {
  if (!filename || !points || !cells || n_points < 0 || n_cells < 0 || cell_type < IU_TRIANGLE_CT || cell_type > IU_TETRA_CT) {
    set_error("iu_b200_write_vtk_arrays: bad argument");
    return IU_B200_EINVAL;
  }
  auto cols = [](int n, const void* data, const char* const* names, const char* prefix) {
    VtuColumns c;
    c.data = data;
    for (int v = 0; v < n; ++v) c.names.push_back(names && names[v] && names[v][0] ? names[v] : prefix + std::to_string(v + 1));
    return c;
  };
  return write_vtu(filename, cell_type, n_points, n_cells, points, cells, cols(n_cell_data, cell_data, cell_names, "cell_data_"),
                   cols(n_icell_data, icell_data, icell_names, "icell_data_"), cols(n_point_data, point_data, point_names, "point_data_"))
             ? IU_B200_OK : IU_B200_EIO;
}

int iu_b200_build_cell_neighbors(int cell_type, int32_t n_points, int32_t n_cells, const double* points,
                                 const int32_t* cells, int32_t* neighbors) {
  if (!points || !cells || !neighbors || n_points < 0 || n_cells < 0) { set_error("bad argument"); return IU_B200_EINVAL; }
  try {
    return build_cell_neighbors(cell_type, n_points, n_cells, points, cells, neighbors) ? IU_B200_OK : IU_B200_EINVAL;
  } catch (const std::bad_alloc&) {
    set_error("out of host memory");
    return IU_B200_ENOMEM;
  }
}

void* iu_b200_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) {
    set_error("cudaMallocHost failed");
    return nullptr;
  }
  return p;
}
void iu_b200_host_free(void* p) {
  if (p) cudaFreeHost(p);
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// field-line tracer (iu_integrate_along_field)
// ------------------------------------------------------------------------------------------------
namespace {

int trace_args_ok(const iu_b200_grid* g, int64_t n, int ndim, int nvar, int integrand, const int32_t* i_field, double min_dx,
                  double max_dx, int max_steps, const void* y, const void* n_steps, int mask) {
  if (!g) { set_error("null grid"); return IU_B200_EINVAL; }
  if (n < 0) { set_error("negative count"); return IU_B200_EINVAL; }
  if (ndim < 1 || ndim > 3 || nvar < 0 || nvar > TRACE_MAX_NVAR || !i_field) { set_error("need ndim in 1..3, nvar in 0..3"); return IU_B200_EINVAL; }
  if (integrand < 0 || integrand > 2 || (integrand == 0 && nvar != 0)) { set_error("unknown integrand (0 needs nvar = 0)"); return IU_B200_EINVAL; }
  for (int k = 0; k < ndim; ++k)
    if (i_field[k] < 1 || i_field[k] > g->dg.n_pd) { set_error("i_field: column outside 1..n_point_data"); return IU_B200_EINVAL; }
  if (max_dx < min_dx) { set_error("max_dx < min_dx"); return IU_B200_EINVAL; }              // :1036
  if (max_steps < 1) { set_error("max_steps < 1"); return IU_B200_EINVAL; }                  // :1037
  if (mask < 0 || mask > g->dg.n_icd) { set_error("i_icell_mask outside 0..n_icell_data"); return IU_B200_EINVAL; }
  if (n > 0 && (!y || !n_steps)) { set_error("null array"); return IU_B200_EINVAL; }
  return IU_B200_OK;
}

// y, y_field, n_steps, boundary_material are device pointers
void run_trace(iu_b200_grid* g, int64_t n, int ndim, int nvar, int integrand, const int32_t* i_field, double min_dx, double max_dx,
               int max_steps, double rtol, double atol, int reverse, double* y, double* y_field, int32_t* n_steps,
               int axisymmetric, int i_icell_mask, int mask_value, int32_t* boundary_material, cudaStream_t s) {
  WorkspaceScope ws(g, s);
  g->tr_seed.ensure((size_t)3 * n);
  g->tr_field.ensure((size_t)ndim * n);
  g->tr_cell.ensure((size_t)n);
  g->n_launches += launch_trace_seeds(n, ndim, ndim + nvar, max_steps, y, g->tr_seed.p, s);
  CU(cudaMemsetAsync(g->tr_cell.p, 0, (size_t)n * 4, s));                                    // i_cell = 0: cold, :1051
  run_query(g, make_args(g, QM_INTERP, n, g->tr_seed.p, ndim, i_field, 0, g->tr_cell.p, g->tr_field.p, nullptr), s);
  TraceArgs a;
  a.n = n; a.ndim = ndim; a.nvar = nvar; a.integrand = integrand;
  for (int k = 0; k < ndim; ++k) a.ivars[k] = i_field[k] - 1;
  a.min_dx = min_dx; a.max_dx = max_dx; a.rtol = rtol; a.atol = atol; a.max_steps = max_steps;
  a.reverse = reverse != 0; a.axisymmetric = axisymmetric != 0;
  a.mask_col = i_icell_mask - 1; a.mask_value = mask_value; a.max_walk_steps = (int)g->max_walk_steps;
  a.y = y; a.y_field = y_field; a.n_steps = n_steps; a.boundary_material = boundary_material;
  a.i_cell0 = g->tr_cell.p; a.field0 = g->tr_field.p;
  if (want_sort(g, QM_CELL, n) && n < (1LL << 31)) {   // lines in Morton order of their seeds
    const int end_bit = morton_end_bit(g->dg);
    g->keys.ensure(n); g->keys_alt.ensure(n); g->idx.ensure(n); g->idx_alt.ensure(n);
    g->sort_temp.ensure(sort_temp_bytes(n, end_bit));
    g->n_launches += launch_morton_sort(g->dg, n, g->tr_seed.p, nullptr, g->keys.p, g->keys_alt.p, g->idx.p, g->idx_alt.p,
                                        g->sort_temp.p, g->sort_temp.n, nullptr, s);
    a.perm = g->idx_alt.p;
  }
  g->n_launches += launch_trace(g->dg, a, s);
  CU(cudaGetLastError());
}

}  // namespace

extern "C" {

int iu_b200_integrate_along_field_batch_dev(iu_b200_grid* g, int64_t n, int ndim, int nvar, int integrand, const int32_t* i_field,
                                            double min_dx, double max_dx, int max_steps, double rtol, double atol, int reverse,
                                            double* y, double* y_field, int32_t* n_steps, int axisymmetric, int i_icell_mask,
                                            int mask_value, int32_t* boundary_material, void* stream) {
  int rc = trace_args_ok(g, n, ndim, nvar, integrand, i_field, min_dx, max_dx, max_steps, y, n_steps, i_icell_mask);
  if (rc != IU_B200_OK || n == 0) return rc;
  std::lock_guard<std::mutex> lock(g->mu);
  return guarded([&]() {
    DeviceGuard guard(g->device);
    run_trace(g, n, ndim, nvar, integrand, i_field, min_dx, max_dx, max_steps, rtol, atol, reverse, y, y_field, n_steps,
              axisymmetric, i_icell_mask, mask_value, boundary_material, (cudaStream_t)stream);
    return IU_B200_OK;
  });
}

// Host buffers: field lines are processed in chunks whose curves fit a bounded device buffer;
// y / y_field go up first because columns beyond n_steps must come back untouched.
static int trace_host_one(iu_b200_grid* g, int64_t n, int ndim, int nvar, int integrand, const int32_t* i_field,
                          double min_dx, double max_dx, int max_steps, double rtol, double atol, int reverse,
                          double* y, double* y_field, int32_t* n_steps, int axisymmetric, int i_icell_mask,
                          int mask_value, int32_t* boundary_material);
int iu_b200_integrate_along_field_batch(iu_b200_grid* g, int64_t n, int ndim, int nvar, int integrand, const int32_t* i_field,
                                        double min_dx, double max_dx, int max_steps, double rtol, double atol, int reverse,
                                        double* y, double* y_field, int32_t* n_steps, int axisymmetric, int i_icell_mask,
                                        int mask_value, int32_t* boundary_material) {
  int rc = trace_args_ok(g, n, ndim, nvar, integrand, i_field, min_dx, max_dx, max_steps, y, n_steps, i_icell_mask);
  if (rc != IU_B200_OK || n == 0) return rc;
  const size_t ny = (size_t)(ndim + nvar) * max_steps, nf = (size_t)ndim * max_steps;
  return run_slabs(g, n, [&](iu_b200_grid* gd, int64_t off, int64_t m) {
    return trace_host_one(gd, m, ndim, nvar, integrand, i_field, min_dx, max_dx, max_steps, rtol, atol, reverse, y + ny * off,
                          y_field ? y_field + nf * off : nullptr, n_steps + off, axisymmetric, i_icell_mask, mask_value,
                          boundary_material ? boundary_material + off : nullptr);
  }, 4096);
}
static int trace_host_one(iu_b200_grid* g, int64_t n, int ndim, int nvar, int integrand, const int32_t* i_field,
                          double min_dx, double max_dx, int max_steps, double rtol, double atol, int reverse,
                          double* y, double* y_field, int32_t* n_steps, int axisymmetric, int i_icell_mask,
                          int mask_value, int32_t* boundary_material) {
  int rc = trace_args_ok(g, n, ndim, nvar, integrand, i_field, min_dx, max_dx, max_steps, y, n_steps, i_icell_mask);
  if (rc != IU_B200_OK || n == 0) return rc;
  std::lock_guard<std::mutex> lock(g->mu);
  return guarded([&]() {
    DeviceGuard guard(g->device);
    const size_t ny = (size_t)(ndim + nvar) * max_steps, nf = (size_t)ndim * max_steps;
    const int64_t chunk = std::max<int64_t>(1, std::min<int64_t>(n, (int64_t)((size_t(1) << 28) / (ny * 8))));   // <= 256 MB of curves
    DevBuf<double> d_y, d_f;
    DevBuf<int32_t> d_ns, d_bm;
    d_y.alloc(ny * chunk);
    if (y_field) d_f.alloc(nf * chunk);
    d_ns.alloc(chunk);
    if (boundary_material) d_bm.alloc(chunk);
    cudaStream_t s = g->s_comp;
    for (int64_t off = 0; off < n; off += chunk) {
      const int64_t m = std::min(chunk, n - off);
      CU(cudaMemcpyAsync(d_y.p, y + ny * off, ny * m * 8, cudaMemcpyHostToDevice, s));
      if (y_field) CU(cudaMemcpyAsync(d_f.p, y_field + nf * off, nf * m * 8, cudaMemcpyHostToDevice, s));
      if (boundary_material) CU(cudaMemcpyAsync(d_bm.p, boundary_material + off, (size_t)m * 4, cudaMemcpyHostToDevice, s));
      run_trace(g, m, ndim, nvar, integrand, i_field, min_dx, max_dx, max_steps, rtol, atol, reverse, d_y.p, d_f.p, d_ns.p,
                axisymmetric, i_icell_mask, mask_value, d_bm.p, s);
      CU(cudaMemcpyAsync(y + ny * off, d_y.p, ny * m * 8, cudaMemcpyDeviceToHost, s));
      if (y_field) CU(cudaMemcpyAsync(y_field + nf * off, d_f.p, nf * m * 8, cudaMemcpyDeviceToHost, s));
      CU(cudaMemcpyAsync(n_steps + off, d_ns.p, (size_t)m * 4, cudaMemcpyDeviceToHost, s));
      if (boundary_material) CU(cudaMemcpyAsync(boundary_material + off, d_bm.p, (size_t)m * 4, cudaMemcpyDeviceToHost, s));
      CU(cudaStreamSynchronize(s));
    }
    return IU_B200_OK;
  });
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// the hot path on host buffers: three-slot pipeline  H2D (s_in) | kernels (s_comp) | D2H (s_out)
// ------------------------------------------------------------------------------------------------
namespace {

void* stage_buffer(iu_b200_grid* g, size_t bytes) {
  if (bytes > g->h_stage_bytes) {
    if (g->h_stage) cudaFreeHost(g->h_stage);
    g->h_stage = nullptr;
    g->h_stage_bytes = 0;
    CU(cudaMallocHost(&g->h_stage, bytes));
    g->h_stage_bytes = bytes;
  }
  return g->h_stage;
}

// H2D(slot, off, m) enqueues on s_in; COMPUTE(slot, m) on s_comp; D2H(slot, off, m, merge) on s_out.
// `merge` is true when the chunk's `trigger` counter moved, i.e. some output rows must stay
// untouched in the caller's arrays; D2H then goes through pinned staging and a host-side row copy.
template <class H2D, class COMPUTE, class D2H>
int run_pipeline(iu_b200_grid* g, int64_t n, int trigger, H2D&& h2d, COMPUTE&& compute, D2H&& d2h) {
  CU(cudaDeviceSynchronize());
  unsigned long long last[CNT_N], base[CNT_N];
  CU(cudaMemcpy(base, g->dg.counters, sizeof(base), cudaMemcpyDeviceToHost));
  std::memcpy(last, base, sizeof(last));
  const int64_t chunk = g->host_chunk;
  const int64_t n_chunks = (n + chunk - 1) / chunk;
  for (int64_t c = 0; c <= n_chunks; ++c) {
    if (c < n_chunks) {
      Slot& sl = g->slot[c % N_SLOTS];
      const int64_t off = c * chunk, m = std::min(chunk, n - off);
      if (c >= N_SLOTS) CU(cudaStreamWaitEvent(g->s_in, sl.out_done, 0));
      h2d(sl, off, m);
      CU(cudaEventRecord(sl.in_done, g->s_in));
      CU(cudaStreamWaitEvent(g->s_comp, sl.in_done, 0));
      compute(sl, m);
      CU(cudaMemcpyAsync(g->h_counters + (c % N_SLOTS) * CNT_N, g->dg.counters, CNT_N * sizeof(unsigned long long),
                         cudaMemcpyDeviceToHost, g->s_comp));
      CU(cudaEventRecord(sl.comp_done, g->s_comp));
    }
    const int64_t d = c - 1;
    if (d >= 0) {
      Slot& sl = g->slot[d % N_SLOTS];
      const int64_t off = d * chunk, m = std::min(chunk, n - off);
      CU(cudaEventSynchronize(sl.comp_done));
      const unsigned long long* now = g->h_counters + (d % N_SLOTS) * CNT_N;
      const bool merge = now[trigger] != last[trigger];
      std::memcpy(last, now, sizeof(last));
      CU(cudaStreamWaitEvent(g->s_out, sl.comp_done, 0));
      d2h(sl, off, m, merge);
      CU(cudaEventRecord(sl.out_done, g->s_out));
    }
  }
  CU(cudaStreamSynchronize(g->s_out));
  const bool bad = last[CNT_HINT_ERROR] != base[CNT_HINT_ERROR];
  g->seen[CNT_HINT_ERROR] = last[CNT_HINT_ERROR];
  if (bad) {
    set_error("i_cell > ug%n_cells (or a start cell outside 1..n_cells)");
    return IU_B200_EHINT;
  }
  return IU_B200_OK;
}

int query_host_one(iu_b200_grid* g, int mode, int64_t n, const double* r, int n_vars, const int32_t* i_vars, int i_var,
                   int32_t* i_cell, double* out_f64, int32_t* out_i32, double* out2_f64) {
  int rc = check_query_args(g, n, r, mode, n_vars, i_vars, i_var);
  if (rc != IU_B200_OK || n == 0) return rc;
  if (mode != QM_NEAREST && !i_cell) { set_error("null i_cell array"); return IU_B200_EINVAL; }
  const bool fused = mode == QM_INTERP_CELL;
  const bool f64_out = mode == QM_INTERP || mode == QM_INTERP_ICELL || mode == QM_CELL_F64 || fused;
  const bool i32_out = mode == QM_NEAREST || mode == QM_CELL_I32;
  if ((f64_out && !out_f64) || (i32_out && !out_i32) || (fused && !out2_f64)) { set_error("null output array"); return IU_B200_EINVAL; }
  const size_t row = mode == QM_INTERP || mode == QM_INTERP_ICELL || fused ? (size_t)n_vars : 1;   // output words per query
  std::lock_guard<std::mutex> lock(g->mu);
  return guarded([&]() {
    DeviceGuard guard(g->device);
    const size_t cap = (size_t)std::min<int64_t>(g->host_chunk, n);
    for (Slot& sl : g->slot) {
      sl.r.ensure(3 * cap);
      if (mode != QM_NEAREST) sl.i_a.ensure(cap);
      if (f64_out) sl.out_f64.ensure(row * cap);
      if (i32_out) sl.i_b.ensure(cap);
      if (fused) sl.r1.ensure(cap);
    }
    auto h2d = [&](Slot& sl, int64_t off, int64_t m) {
      CU(cudaMemcpyAsync(sl.r.p, r + 3 * off, (size_t)m * 24, cudaMemcpyHostToDevice, g->s_in));
      if (mode == QM_NEAREST) return;
      // a chunk without a single hint (benchmark.f90's first pass: i_cell(:) = 0) does not send its i_cell over the bus:
      // the host scans the 4 bytes per query (while the previous chunks are in flight) and the device buffer is zeroed
      const int32_t* h = i_cell + off;
      int any_hint = 0;
      for (int64_t i = 0; i < m; ++i) any_hint |= h[i] > 0;
      if (any_hint) CU(cudaMemcpyAsync(sl.i_a.p, h, (size_t)m * 4, cudaMemcpyHostToDevice, g->s_in));
      else { CU(cudaMemsetAsync(sl.i_a.p, 0, (size_t)m * 4, g->s_in)); g->h2d_skipped_bytes += (size_t)m * 4; }
    };
    auto compute = [&](Slot& sl, int64_t m) {
      run_query(g, make_args(g, mode, m, sl.r.p, n_vars, i_vars, i_var, sl.i_a.p, sl.out_f64.p, sl.i_b.p, fused ? sl.r1.p : nullptr), g->s_comp);
    };
    auto d2h = [&](Slot& sl, int64_t off, int64_t m, bool merge) {
      if (mode != QM_NEAREST && mode != QM_INTERP_ICELL)
        CU(cudaMemcpyAsync(i_cell + off, sl.i_a.p, (size_t)m * 4, cudaMemcpyDeviceToHost, g->s_out));
      if (mode == QM_CELL) return;
      const size_t row_bytes = f64_out ? row * 8 : 4;
      char* dst = f64_out ? (char*)(out_f64 + row * off) : (char*)(out_i32 + off);
      const void* src = f64_out ? (const void*)sl.out_f64.p : (const void*)sl.i_b.p;
      if (!merge || mode == QM_NEAREST) {
        CU(cudaMemcpyAsync(dst, src, (size_t)m * row_bytes, cudaMemcpyDeviceToHost, g->s_out));
        if (fused) CU(cudaMemcpyAsync(out2_f64 + off, sl.r1.p, (size_t)m * 8, cudaMemcpyDeviceToHost, g->s_out));
        return;
      }
      // some queries of this chunk were not found: their rows stay untouched (:447, :461, :492)
      char* st = (char*)stage_buffer(g, (size_t)g->host_chunk * (row_bytes + (fused ? 8 : 0)));
      char* st2 = st + (size_t)g->host_chunk * row_bytes;
      CU(cudaMemcpyAsync(st, src, (size_t)m * row_bytes, cudaMemcpyDeviceToHost, g->s_out));
      if (fused) CU(cudaMemcpyAsync(st2, sl.r1.p, (size_t)m * 8, cudaMemcpyDeviceToHost, g->s_out));
      CU(cudaStreamSynchronize(g->s_out));
      const int32_t* ic = i_cell + off;
      for (int64_t i = 0; i < m; ++i)
        if (ic[i] >= 1) {
          std::memcpy(dst + (size_t)i * row_bytes, st + (size_t)i * row_bytes, row_bytes);
          if (fused) std::memcpy(out2_f64 + off + i, st2 + (size_t)i * 8, 8);
        }
    };
    return run_pipeline(g, n, CNT_NOT_FOUND, h2d, compute, d2h);
  });
}

int query_host(iu_b200_grid* g, int mode, int64_t n, const double* r, int n_vars, const int32_t* i_vars, int i_var,
               int32_t* i_cell, double* out_f64, int32_t* out_i32, double* out2_f64 = nullptr) {
  const size_t row = mode == QM_INTERP || mode == QM_INTERP_ICELL || mode == QM_INTERP_CELL ? (size_t)std::max(n_vars, 0) : 1;
  return run_slabs(g, n, [&](iu_b200_grid* gd, int64_t off, int64_t m) {
    return query_host_one(gd, mode, m, r ? r + 3 * off : nullptr, n_vars, i_vars, i_var, i_cell ? i_cell + off : nullptr,
                          out_f64 ? out_f64 + row * off : nullptr, out_i32 ? out_i32 + off : nullptr,
                          out2_f64 ? out2_f64 + off : nullptr);
  });
}

}  // namespace

extern "C" {

int iu_b200_find_nearby_cell_batch(iu_b200_grid* g, int64_t n, const double* r, int32_t* idx) {
  return query_host(g, QM_NEAREST, n, r, 0, nullptr, 0, nullptr, nullptr, idx);
}
int iu_b200_get_cell_batch(iu_b200_grid* g, int64_t n, const double* r, int32_t* i_cell) {
  return query_host(g, QM_CELL, n, r, 0, nullptr, 0, i_cell, nullptr, nullptr);
}
int iu_b200_interpolate_at_batch(iu_b200_grid* g, int64_t n, const double* r, int n_vars, const int32_t* i_vars,
                                 double* var_at_r, int32_t* i_cell) {
  return query_host(g, QM_INTERP, n, r, n_vars, i_vars, 0, i_cell, var_at_r, nullptr);
}
int iu_b200_interpolate_scalar_at_batch(iu_b200_grid* g, int64_t n, const double* r, int i_var, double* var_at_r,
                                        int32_t* i_cell) {
  const int32_t iv[1] = {i_var};
  return query_host(g, QM_INTERP, n, r, 1, iv, 0, i_cell, var_at_r, nullptr);
}
int iu_b200_interpolate_at_icell_batch(iu_b200_grid* g, int64_t n, const double* r, int n_vars, const int32_t* i_vars,
                                       double* var_at_r, const int32_t* i_cell) {
  return query_host(g, QM_INTERP_ICELL, n, r, n_vars, i_vars, 0, const_cast<int32_t*>(i_cell), var_at_r, nullptr);
}
int iu_b200_interpolate_and_cell_scalar_at_batch(iu_b200_grid* g, int64_t n, const double* r, int n_vars, const int32_t* i_vars,
                                                 int i_cell_var, double* var_at_r, double* cell_res, int32_t* i_cell) {
  return query_host(g, QM_INTERP_CELL, n, r, n_vars, i_vars, i_cell_var, i_cell, var_at_r, nullptr, cell_res);
}
int iu_b200_get_cell_scalar_at_batch(iu_b200_grid* g, int64_t n, const double* r, int i_var, int32_t* i_guess, double* res) {
  return query_host(g, QM_CELL_F64, n, r, 0, nullptr, i_var, i_guess, res, nullptr);
}
int iu_b200_get_icell_scalar_at_batch(iu_b200_grid* g, int64_t n, const double* r, int i_var, int32_t* i_guess, int32_t* res) {
  return query_host(g, QM_CELL_I32, n, r, 0, nullptr, i_var, i_guess, nullptr, res);
}

static int walk_host_one(iu_b200_grid* g, int64_t n, int max_steps, const double* r0, const double* r1, const int32_t* ic0,
                         int32_t* ic1, double* r_p, int32_t* n_steps, int32_t* status, int i_icell_mask);
int iu_b200_get_cell_through_neighbors_batch(iu_b200_grid* g, int64_t n, int max_steps, const double* r0, const double* r1,
                                             const int32_t* ic0, int32_t* ic1, double* r_p, int32_t* n_steps,
                                             int32_t* status, int i_icell_mask) {
  int rc = walk_args_ok(g, n, max_steps, r0, r1, ic0, ic1, r_p, n_steps, status, i_icell_mask);
  if (rc != IU_B200_OK || n == 0) return rc;
  return run_slabs(g, n, [&](iu_b200_grid* gd, int64_t off, int64_t m) {
    return walk_host_one(gd, m, max_steps, r0 + 3 * off, r1 + 3 * off, ic0 + off, ic1 + off, r_p + 3 * off, n_steps + off,
                         status + off, i_icell_mask);
  });
}
static int walk_host_one(iu_b200_grid* g, int64_t n, int max_steps, const double* r0, const double* r1, const int32_t* ic0,
                         int32_t* ic1, double* r_p, int32_t* n_steps, int32_t* status, int i_icell_mask) {
  int rc = walk_args_ok(g, n, max_steps, r0, r1, ic0, ic1, r_p, n_steps, status, i_icell_mask);
  if (rc != IU_B200_OK || n == 0) return rc;
  std::lock_guard<std::mutex> lock(g->mu);
  return guarded([&]() {
    DeviceGuard guard(g->device);
    const size_t cap = (size_t)std::min<int64_t>(g->host_chunk, n);
    for (Slot& sl : g->slot) {
      sl.r.ensure(3 * cap); sl.r1.ensure(3 * cap); sl.out_f64.ensure(3 * cap);
      sl.i_a.ensure(cap); sl.i_b.ensure(cap); sl.i_c.ensure(cap); sl.i_d.ensure(cap);
    }
    auto h2d = [&](Slot& sl, int64_t off, int64_t m) {
      CU(cudaMemcpyAsync(sl.r.p, r0 + 3 * off, (size_t)m * 24, cudaMemcpyHostToDevice, g->s_in));
      CU(cudaMemcpyAsync(sl.r1.p, r1 + 3 * off, (size_t)m * 24, cudaMemcpyHostToDevice, g->s_in));
      CU(cudaMemcpyAsync(sl.i_a.p, ic0 + off, (size_t)m * 4, cudaMemcpyHostToDevice, g->s_in));
    };
    auto compute = [&](Slot& sl, int64_t m) {
      WalkArgs a;
      a.n = m; a.max_steps = std::min(max_steps, MAX_STEPS_CAP); a.r0 = sl.r.p; a.r1 = sl.r1.p; a.ic0 = sl.i_a.p; a.ic1 = sl.i_b.p;
      a.r_p = sl.out_f64.p; a.n_steps = sl.i_c.p; a.status = sl.i_d.p; a.mask_col = i_icell_mask - 1;
      a.mark_early = true;
      run_walk(g, a, g->s_comp);
    };
    auto d2h = [&](Slot& sl, int64_t off, int64_t m, bool merge) {
      CU(cudaMemcpyAsync(ic1 + off, sl.i_b.p, (size_t)m * 4, cudaMemcpyDeviceToHost, g->s_out));
      CU(cudaMemcpyAsync(status + off, sl.i_d.p, (size_t)m * 4, cudaMemcpyDeviceToHost, g->s_out));
      if (!merge) {
        CU(cudaMemcpyAsync(r_p + 3 * off, sl.out_f64.p, (size_t)m * 24, cudaMemcpyDeviceToHost, g->s_out));
        CU(cudaMemcpyAsync(n_steps + off, sl.i_c.p, (size_t)m * 4, cudaMemcpyDeviceToHost, g->s_out));
        return;
      }
      // rays shorter than tiny_distance leave r_p and n_steps untouched (:689-692)
      char* st = (char*)stage_buffer(g, (size_t)g->host_chunk * 28);
      double* st_rp = (double*)st;
      int32_t* st_ns = (int32_t*)(st + (size_t)g->host_chunk * 24);
      CU(cudaMemcpyAsync(st_rp, sl.out_f64.p, (size_t)m * 24, cudaMemcpyDeviceToHost, g->s_out));
      CU(cudaMemcpyAsync(st_ns, sl.i_c.p, (size_t)m * 4, cudaMemcpyDeviceToHost, g->s_out));
      CU(cudaStreamSynchronize(g->s_out));
      for (int64_t i = 0; i < m; ++i)
        if (st_ns[i] != INT32_MIN) {
          n_steps[off + i] = st_ns[i];
          for (int k = 0; k < 3; ++k) r_p[3 * (off + i) + k] = st_rp[3 * i + k];
        }
    };
    return run_pipeline(g, n, CNT_EARLY, h2d, compute, d2h);
  });
}

}  // extern "C"
