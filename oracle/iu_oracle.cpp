// =====================================================================================
// iu_oracle.cpp -- TEST INFRASTRUCTURE ONLY.  NOT PART OF THE PRODUCT.
//
// CPU restatement (C++17, compiled -O2 -ffp-contract=off, never -ffast-math) of the hot
// path of jannisteunissen/interpolate_unstructured (serial Fortran 2008), operation by
// operation in the reference's evaluation order.  Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may load this library; the
// product (libiu_b200.so) never links, imports or calls it.
//
// PARITY STATUS
//   * Derived geometry, ray walk and interpolation follow m_interp_unstructured.f90
//     line by line (citations on every function).  The reference's own tests pin only
//     interpolated values of a linear field (|res-(x+y+z+1)| <= 1e-14 on triangle.vtu,
//     quad.vtu, tetra.vtu; test_triangle.f90:23,39-49 etc.).  tests/test_oracle.py
//     checks exactly those criteria on the reference fixtures.
//   * The reference itself cannot be built here (no Fortran compiler on the image, and
//     the kdtree2 submodule -- github.com/jannisteunissen/kdtree2 @
//     ad2cfe28fa92f5080b737d4ebf31160fd4433347 -- is absent from /root/reference), so
//     there is no oracle/_ref.  The nearest-cell step is therefore "PARITY UNPINNED":
//     KdTree2 below restates M. Kennel's published kdtree2 algorithm (arXiv:physics/0408067;
//     bucket size 12, mean split, near-child-first search, "sd > ballsize -> skip") from
//     the published description; it agrees with brute force (exact fp64 nearest centre)
//     everywhere except on exact distance ties, which are reported so tests can exclude
//     and count them.
// =====================================================================================
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <limits>
#include <memory>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

constexpr int IU_TRIANGLE = 1, IU_QUAD = 2, IU_TETRA = 3;   // m_interp_unstructured.f90:13-15
constexpr int MAXP = 4;                                     // ug_max_points_per_cell, :18
constexpr double TINY_DISTANCE = 1e-100;                    // :20
constexpr double HUGE_DISTANCE = 1e100;                     // :738

thread_local std::string g_err;

struct V3 { double x, y, z; };
inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 neg(V3 a) { return {-a.x, -a.y, -a.z}; }

// Fortran dot_product over 3 elements: 0 + a1*b1 + a2*b2 + a3*b3, left to right.
inline double dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }

// cross_product, m_interp_unstructured.f90:644-651
inline V3 cross(V3 a, V3 b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
// cross_product_z, :654-657
inline double cross_z(V3 a, V3 b) { return a.x * b.y - a.y * b.x; }
// scalar_triple_product, :659-662
inline double stp(V3 a, V3 b, V3 c) { return dot3(a, cross(b, c)); }

// gfortran NORM2 (inline expansion in trans-intrinsic / libgfortran norm2_r8): running-scale
// algorithm, scale starts at 1.  SURVEY.md App. D.
inline double norm2(V3 v) {
  double scale = 1.0, s = 0.0;
  const double a[3] = {v.x, v.y, v.z};
  for (int i = 0; i < 3; ++i) {
    if (a[i] != 0.0) {
      double ax = std::fabs(a[i]);
      if (scale < ax) {
        double val = scale / ax;
        s = 1.0 + s * val * val;
        scale = ax;
      } else {
        double val = ax / scale;
        s += val * val;
      }
    }
  }
  return scale * std::sqrt(s);
}

// ---------------------------------------------------------------------------------
// kdtree2 restatement (see header: published algorithm, source absent => unpinned).
// Call sites in the reference: kdtree2_create(cell_centers, sort=.false.,
// rearrange=.false.) m_interp_unstructured.f90:261; kdtree2_n_nearest(tree, r, 1, res)
// :279; res(1)%idx :280.
// ---------------------------------------------------------------------------------
struct KdTree2 {
  static constexpr int bucket_size = 12;
  struct Node {
    int cut_dim = -1;
    double cut_val = 0, cut_val_left = 0, cut_val_right = 0;
    int l = 0, u = 0;            // inclusive range into ind
    int left = -1, right = -1;   // node indices, -1 = null
    double lo[3], hi[3];         // box
  };
  const double* data = nullptr;  // (3, n) column-major: data[3*i + k]
  int n = 0;
  std::vector<int> ind;
  std::vector<Node> nodes;
  int root = -1;

  double coord(int k, int i) const { return data[3 * (size_t)i + k]; }

  void spread(int c, int l, int u, double& lo, double& hi) const {
    double smin = coord(c, ind[l]), smax = smin;
    for (int i = l + 1; i <= u; ++i) {
      double v = coord(c, ind[i]);
      smin = std::min(smin, v);
      smax = std::max(smax, v);
    }
    lo = smin; hi = smax;
  }

  // select_on_coordinate_value: points <= alpha first; returns index of last point <= alpha.
  int select_on_value(int c, double alpha, int li, int ui) {
    int lb = li, ub = ui;
    while (lb < ub) {
      if (coord(c, ind[lb]) <= alpha) ++lb;
      else { std::swap(ind[lb], ind[ub]); --ub; }
    }
    return (coord(c, ind[lb]) <= alpha) ? lb : lb - 1;
  }

  int build_range(int l, int u, int parent) {
    if (u < l) return -1;
    int me = (int)nodes.size();
    nodes.emplace_back();
    if ((u - l) <= bucket_size) {
      Node nd;
      for (int k = 0; k < 3; ++k) spread(k, l, u, nd.lo[k], nd.hi[k]);
      nd.cut_dim = -1; nd.l = l; nd.u = u;
      nodes[me] = nd;
      return me;
    }
    Node nd;
    for (int k = 0; k < 3; ++k) {
      bool recompute = true;
      if (parent >= 0 && k != nodes[parent].cut_dim) recompute = false;
      if (recompute) spread(k, l, u, nd.lo[k], nd.hi[k]);
      else { nd.lo[k] = nodes[parent].lo[k]; nd.hi[k] = nodes[parent].hi[k]; }
    }
    int c = 0;   // maxloc: first maximum
    for (int k = 1; k < 3; ++k)
      if (nd.hi[k] - nd.lo[k] > nd.hi[c] - nd.lo[c]) c = k;
    double sum = 0.0;
    for (int i = l; i <= u; ++i) sum += coord(c, ind[i]);
    double average = sum / (double)(u - l + 1);
    nd.cut_val = average;
    int m = select_on_value(c, average, l, u);
    nd.cut_dim = c; nd.l = l; nd.u = u;
    nodes[me] = nd;   // children consult parent's approximate box + cut_dim
    int left = build_range(l, m, me);
    int right = build_range(m + 1, u, me);
    Node& r = nodes[me];
    r.left = left; r.right = right;
    if (right < 0) {
      for (int k = 0; k < 3; ++k) { r.lo[k] = nodes[left].lo[k]; r.hi[k] = nodes[left].hi[k]; }
      r.cut_val_left = nodes[left].hi[c];
      r.cut_val = r.cut_val_left;
    } else if (left < 0) {
      for (int k = 0; k < 3; ++k) { r.lo[k] = nodes[right].lo[k]; r.hi[k] = nodes[right].hi[k]; }
      r.cut_val_right = nodes[right].lo[c];
      r.cut_val = r.cut_val_right;
    } else {
      r.cut_val_right = nodes[right].lo[c];
      r.cut_val_left = nodes[left].hi[c];
      r.cut_val = (r.cut_val_left + r.cut_val_right) / 2;
      for (int k = 0; k < 3; ++k) {
        r.hi[k] = std::max(nodes[left].hi[k], nodes[right].hi[k]);
        r.lo[k] = std::min(nodes[left].lo[k], nodes[right].lo[k]);
      }
    }
    return me;
  }

  void create(const double* d, int npts) {
    data = d; n = npts;
    ind.resize(n);
    for (int i = 0; i < n; ++i) ind[i] = i;
    nodes.clear();
    nodes.reserve((size_t)n / 4 + 16);
    root = build_range(0, n - 1, -1);
  }

  struct Search { const double* qv; double ballsize; int nfound; int idx; };

  static double dis2_from_bnd(double x, double amin, double amax) {
    if (x > amax) return (x - amax) * (x - amax);
    if (x < amin) return (amin - x) * (amin - x);
    return 0.0;
  }

  void process_terminal(const Node& nd, Search& sr) const {
    for (int i = nd.l; i <= nd.u; ++i) {
      int id = ind[i];
      double sd = 0.0;
      bool skip = false;
      for (int k = 0; k < 3; ++k) {
        double t = coord(k, id) - sr.qv[k];
        sd = sd + t * t;
        if (sd > sr.ballsize) { skip = true; break; }
      }
      if (skip) continue;
      // nn = 1: first hit fills the list, later hits (sd <= ballsize) replace the max.
      sr.nfound = 1; sr.ballsize = sd; sr.idx = id;
    }
  }

  void search(int ni, Search& sr) const {
    const Node& nd = nodes[ni];
    if (nd.left < 0 && nd.right < 0) { process_terminal(nd, sr); return; }
    int c = nd.cut_dim;
    double qval = sr.qv[c];
    int ncloser, nfarther; double dis;
    if (qval < nd.cut_val) { ncloser = nd.left; nfarther = nd.right; dis = (nd.cut_val_right - qval) * (nd.cut_val_right - qval); }
    else { ncloser = nd.right; nfarther = nd.left; dis = (nd.cut_val_left - qval) * (nd.cut_val_left - qval); }
    if (ncloser >= 0) search(ncloser, sr);
    if (nfarther >= 0) {
      double ballsize = sr.ballsize;
      if (dis <= ballsize) {
        for (int k = 0; k < 3; ++k) {
          if (k != c) {
            dis += dis2_from_bnd(sr.qv[k], nd.lo[k], nd.hi[k]);
            if (dis > ballsize) return;
          }
        }
        search(nfarther, sr);
      }
    }
  }

  // returns 0-based index of the nearest point
  int nearest(const double* qv) const {
    Search sr{qv, (double)std::numeric_limits<float>::max(), 0, -1};   // ballsize = huge(1.0)
    search(root, sr);
    return sr.idx;
  }
};

// ---------------------------------------------------------------------------------
// iu_grid_t, m_interp_unstructured.f90:24-59.  Fortran column-major arrays are stored
// flat with the first index fastest; all indices kept 1-based as in the reference.
// ---------------------------------------------------------------------------------
struct Grid {
  int n_cells = 0, n_points = 0, ppc = 0, cell_type = 0;
  int n_point_data = 0, n_cell_data = 0, n_icell_data = 0;
  std::vector<double> points;          // (3, Np)
  std::vector<int> cells;              // (ppc, Nc) 1-based
  std::vector<int> neighbors;          // (ppc, Nc) 1-based, <1 = none
  std::vector<double> cell_points;     // (3, ppc, Nc)
  std::vector<double> cell_volume;     // (Nc)
  std::vector<double> normals;         // (3, ppc, Nc)
  std::vector<uint8_t> point_is_at_boundary;
  std::vector<double> point_data;      // (Np, nvar)
  std::vector<double> cell_data;       // (Nc, nvar)
  std::vector<int> icell_data;         // (Nc, nvar)
  std::vector<std::string> point_names, cell_names, icell_names;
  std::vector<double> centers;         // (3, Nc)
  KdTree2 tree;

  V3 cp(int k, int n) const {   // cell_points(:, k, n), k and n 1-based
    const double* p = &cell_points[3 * ((size_t)(n - 1) * ppc + (k - 1))];
    return {p[0], p[1], p[2]};
  }
  V3 nrm(int k, int n) const {
    const double* p = &normals[3 * ((size_t)(n - 1) * ppc + (k - 1))];
    return {p[0], p[1], p[2]};
  }
  int nb(int k, int n) const { return neighbors[(size_t)(n - 1) * ppc + (k - 1)]; }
  int cell(int k, int n) const { return cells[(size_t)(n - 1) * ppc + (k - 1)]; }
};

// sum(cell_points(:,:,n), dim=2) / ppc  -- :268, :319, :345, :429
V3 cell_center(const Grid& g, int n) {
  V3 s = g.cp(1, n);
  for (int k = 2; k <= g.ppc; ++k) s = s + g.cp(k, n);
  double d = (double)g.ppc;
  return {s.x / d, s.y / d, s.z / d};
}

// set_cell_points, :291-302
void set_cell_points(Grid& g) {
  g.cell_points.resize((size_t)3 * g.ppc * g.n_cells);
  for (int n = 1; n <= g.n_cells; ++n)
    for (int k = 1; k <= g.ppc; ++k) {
      int p = g.cell(k, n);
      for (int d = 0; d < 3; ++d)
        g.cell_points[3 * ((size_t)(n - 1) * g.ppc + (k - 1)) + d] = g.points[3 * (size_t)(p - 1) + d];
    }
}

// set_face_normal_vectors, :306-370
void set_face_normal_vectors(Grid& g) {
  g.normals.resize((size_t)3 * g.ppc * g.n_cells);
  g.point_is_at_boundary.assign(g.n_points, 0);
  const int ppc = g.ppc;
  for (int n = 1; n <= g.n_cells; ++n) {
    V3 center = cell_center(g, n);
    V3 normal_cell{0, 0, 0};
    if (g.cell_type != IU_TETRA)
      normal_cell = cross(g.cp(2, n) - g.cp(1, n), g.cp(3, n) - g.cp(2, n));   // :322-324
    for (int k = 1; k <= ppc; ++k) {
      int k1 = k % ppc + 1;
      int k2 = (k + 1) % ppc + 1;
      V3 nf;
      if (g.cell_type != IU_TETRA) {
        V3 vec = g.cp(k1, n) - g.cp(k, n);
        nf = cross(vec, normal_cell);                                           // :329-330
      } else {
        nf = cross(g.cp(k1, n) - g.cp(k, n), g.cp(k2, n) - g.cp(k1, n));        // :351-353
      }
      V3 vec = g.cp(k, n) - center;
      if (dot3(vec, nf) < 0) nf = neg(nf);                                      // :334, :357
      double len = norm2(nf);
      double* out = &g.normals[3 * ((size_t)(n - 1) * ppc + (k - 1))];
      out[0] = nf.x / len; out[1] = nf.y / len; out[2] = nf.z / len;            // :336, :359
      if (g.nb(k, n) < 1) {
        g.point_is_at_boundary[g.cell(k, n) - 1] = 1;
        g.point_is_at_boundary[g.cell(k1, n) - 1] = 1;
        if (g.cell_type == IU_TETRA) g.point_is_at_boundary[g.cell(k2, n) - 1] = 1;
      }
    }
  }
}

// set_cell_volumes, :372-410
void set_cell_volumes(Grid& g) {
  g.cell_volume.resize(g.n_cells);
  const double sixth = 1 / 6.0;
  for (int n = 1; n <= g.n_cells; ++n) {
    V3 p1 = g.cp(1, n), p2 = g.cp(2, n), p3 = g.cp(3, n);
    if (g.cell_type == IU_TRIANGLE) {
      g.cell_volume[n - 1] = 0.5 * norm2(cross(p2 - p1, p3 - p1));
    } else if (g.cell_type == IU_QUAD) {
      V3 p4 = g.cp(4, n);
      double a1 = 0.5 * norm2(cross(p2 - p1, p3 - p1));
      double a2 = 0.5 * norm2(cross(p3 - p1, p4 - p1));
      g.cell_volume[n - 1] = a1 + a2;
    } else {
      V3 p4 = g.cp(4, n);
      g.cell_volume[n - 1] = sixth * stp(p2 - p1, p3 - p1, p4 - p1);
    }
  }
}

// build_kdtree, :251-262
void build_kdtree(Grid& g) {
  g.centers.resize((size_t)3 * g.n_cells);
  for (int n = 1; n <= g.n_cells; ++n) {
    V3 c = cell_center(g, n);
    g.centers[3 * (size_t)(n - 1)] = c.x;
    g.centers[3 * (size_t)(n - 1) + 1] = c.y;
    g.centers[3 * (size_t)(n - 1) + 2] = c.z;
  }
  g.tree.create(g.centers.data(), g.n_cells);
}

void finish_grid(Grid& g) {   // tail of iu_read_grid, :915-925
  set_cell_points(g);
  set_face_normal_vectors(g);
  set_cell_volumes(g);
  build_kdtree(g);
}

// find_nearby_cell_kdtree, :272-288 (returns 1-based)
int find_nearby_cell_kdtree(const Grid& g, const double* r) { return g.tree.nearest(r) + 1; }

// exact brute force: strict '<' keeps the lowest index among exact ties.  Also reports
// best and second-best squared distances (second over OTHER indices; equal => tie).
int find_nearby_cell_brute(const Grid& g, const double* r, double* best_out, double* second_out) {
  double best = std::numeric_limits<double>::infinity(), second = best;
  int bi = 0;
  for (int i = 0; i < g.n_cells; ++i) {
    const double* c = &g.centers[3 * (size_t)i];
    double sd = 0.0;
    for (int k = 0; k < 3; ++k) { double t = c[k] - r[k]; sd = sd + t * t; }
    if (sd < best) { second = best; best = sd; bi = i + 1; }
    else if (sd < second) second = sd;
  }
  if (best_out) *best_out = best;
  if (second_out) *second_out = second;
  return bi;
}

struct WalkDiag { double min_abs_left; double min_face_gap; };

// get_cell_intersection, :729-764
void get_cell_intersection(const Grid& g, V3 u, V3& r_p, int ic, double& face_distance, int& i_face,
                           WalkDiag* diag) {
  face_distance = HUGE_DISTANCE;
  i_face = -1;
  double second = HUGE_DISTANCE;
  for (int k = 1; k <= g.ppc; ++k) {
    V3 fn = g.nrm(k, ic);
    double pdn = dot3(u, fn);
    if (pdn > 0) {
      V3 rf = g.cp(k, ic);
      double dist = dot3(rf - r_p, fn) / pdn;
      if (dist < face_distance) { second = face_distance; face_distance = dist; i_face = k; }
      else if (dist < second) second = dist;
    }
  }
  if (diag && second < HUGE_DISTANCE) diag->min_face_gap = std::min(diag->min_face_gap, second - face_distance);
  r_p = {r_p.x + face_distance * u.x, r_p.y + face_distance * u.y, r_p.z + face_distance * u.z};
}

// iu_get_cell_through_neighbors, :667-725.  i_mask = 0 => optional argument absent.
void walk(const Grid& g, int max_steps, V3 r0, V3 r1, int ic0, int& ic1, V3& r_p, int& n_steps,
          int& status, int i_mask, WalkDiag* diag) {
  V3 dvec = r1 - r0;
  double distance_left = norm2(dvec);
  status = 0;
  if (distance_left < TINY_DISTANCE) { ic1 = ic0; return; }   // r_p, n_steps left undefined (:689-692)
  V3 u{dvec.x / distance_left, dvec.y / distance_left, dvec.z / distance_left};
  r_p = r0;
  ic1 = ic0;
  // Fortran "do n_steps = 1, max_steps": after normal loop exhaustion n_steps = max_steps+1.
  for (n_steps = 1; n_steps <= max_steps; ++n_steps) {
    double fd; int i_face;
    get_cell_intersection(g, u, r_p, ic1, fd, i_face, diag);
    distance_left = distance_left - fd;
    if (diag) diag->min_abs_left = std::min(diag->min_abs_left, std::fabs(distance_left));
    if (distance_left > 0) {
      ic1 = g.nb(i_face, ic1);
      if (ic1 < 1) { status = -1; break; }
      if (i_mask > 0) {
        if (g.icell_data[(size_t)(i_mask - 1) * g.n_cells + (ic0 - 1)] !=
            g.icell_data[(size_t)(i_mask - 1) * g.n_cells + (ic1 - 1)]) { status = 1; break; }
      }
    } else {
      break;
    }
  }
}

// iu_get_cell, :414-434.  nn_mode 0 = kdtree2 restatement, 1 = brute force.
void get_cell(const Grid& g, const double* r, int& i_cell, int nn_mode, int* n_steps_out, WalkDiag* diag) {
  int i_start;
  if (i_cell < 1) i_start = nn_mode == 0 ? find_nearby_cell_kdtree(g, r) : find_nearby_cell_brute(g, r, nullptr, nullptr);
  else i_start = i_cell;
  // "Cell index is out of range" error stop of find_nearby_cell_kdtree (:282-286): reachable only when no
  // centre has a comparable distance (infinite / NaN coordinates); reported as "not found" instead of aborting
  if (i_start < 1 || i_start > g.n_cells) { i_cell = 0; if (n_steps_out) *n_steps_out = 0; return; }
  V3 r0 = cell_center(g, i_start);
  V3 r_p{0, 0, 0};
  int n_steps = 0, status = 0;
  walk(g, std::numeric_limits<int>::max() - 1, r0, {r[0], r[1], r[2]}, i_start, i_cell, r_p, n_steps, status, 0, diag);
  if (n_steps_out) *n_steps_out = n_steps;
}

// interpolate_triangle, :529-551
void interpolate_triangle(int nv, const V3* p, const double (*pd)[MAXP], double area, V3 r, double* res) {
  double a[3];
  a[0] = 0.5 * norm2(cross(r - p[1], r - p[2]));
  a[1] = 0.5 * norm2(cross(r - p[2], r - p[0]));
  a[2] = 0.5 * norm2(cross(r - p[0], r - p[1]));
  double inv_area = 1 / area;
  for (int n = 0; n < nv; ++n) res[n] = ((pd[n][0] * a[0] + pd[n][1] * a[1]) + pd[n][2] * a[2]) * inv_area;
}

// interpolate_tetrahedron, :555-586
void interpolate_tetrahedron(int nv, const V3* p, const double (*pd)[MAXP], double vol, V3 r, double* res) {
  V3 v1r = r - p[0], v2r = r - p[1];
  V3 v12 = p[1] - p[0], v13 = p[2] - p[0], v14 = p[3] - p[0], v23 = p[2] - p[1], v24 = p[3] - p[1];
  double w[4];
  w[0] = stp(v2r, v24, v23);
  w[1] = stp(v1r, v13, v14);
  w[2] = stp(v1r, v14, v12);
  w[3] = stp(v1r, v12, v13);
  double inv_weight = 1 / (6 * vol);
  for (int n = 0; n < nv; ++n)
    res[n] = (((pd[n][0] * w[0] + pd[n][1] * w[1]) + pd[n][2] * w[2]) + pd[n][3] * w[3]) * inv_weight;
}

// interpolate_quad, :591-641
void interpolate_quad(int nv, const V3* p, const double (*pd)[MAXP], V3 r, double* res) {
  const double tiny_value = 1e-20;
  V3 q = r - p[0];
  V3 b1 = p[1] - p[0];
  V3 b2 = p[3] - p[0];
  V3 b3 = ((p[0] - p[1]) - p[3]) + p[2];
  double A = cross_z(b2, b3);
  double B = cross_z(b3, q) - cross_z(b1, b2);
  double C = cross_z(b1, q);
  double discrim = B * B - 4 * A * C;
  double c2;
  if (std::fabs(A) < tiny_value) c2 = -C / B;
  else c2 = 0.5 * (-B - std::sqrt(discrim)) / A;
  double den[3] = {b1.x + c2 * b3.x, b1.y + c2 * b3.y, b1.z + c2 * b3.z};
  int dim = 0;   // maxloc(abs(denom)): first maximum
  for (int k = 1; k < 3; ++k) if (std::fabs(den[k]) > std::fabs(den[dim])) dim = k;
  const double qa[3] = {q.x, q.y, q.z}, b2a[3] = {b2.x, b2.y, b2.z};
  double c1 = (qa[dim] - b2a[dim] * c2) / den[dim];
  for (int n = 0; n < nv; ++n) {
    double t1 = pd[n][0] * (1 - c1) + pd[n][1] * c1;
    double t2 = pd[n][3] * (1 - c1) + pd[n][2] * c1;
    res[n] = t1 * (1 - c2) + t2 * c2;
  }
}

// iu_interpolate_at_icell, :498-527
void interpolate_at_icell(const Grid& g, const double* r, int nv, const int* i_vars, double* out, int ic) {
  constexpr int CH = 16;
  V3 p[MAXP];
  for (int k = 1; k <= g.ppc; ++k) p[k - 1] = g.cp(k, ic);
  V3 rr{r[0], r[1], r[2]};
  for (int v0 = 0; v0 < nv; v0 += CH) {
    int m = std::min(CH, nv - v0);
    double pd[CH][MAXP];
    for (int n = 0; n < m; ++n)
      for (int k = 1; k <= g.ppc; ++k)
        pd[n][k - 1] = g.point_data[(size_t)(i_vars[v0 + n] - 1) * g.n_points + (g.cell(k, ic) - 1)];
    if (g.cell_type == IU_TRIANGLE) interpolate_triangle(m, p, pd, g.cell_volume[ic - 1], rr, out + v0);
    else if (g.cell_type == IU_QUAD) interpolate_quad(m, p, pd, rr, out + v0);
    else interpolate_tetrahedron(m, p, pd, g.cell_volume[ic - 1], rr, out + v0);
  }
}

// iu_point_is_inside_cell, :767-786
bool point_is_inside_cell(const Grid& g, const double* r, int ic) {
  const double small_number = 1e-10;
  V3 rr{r[0], r[1], r[2]};
  for (int k = 1; k <= g.ppc; ++k)
    if (dot3(g.cp(k, ic) - rr, g.nrm(k, ic)) < -small_number) return false;
  return true;
}

// iu_integrate_along_field, :989-1217 (Bogacki-Shampine RK3(2) along the interpolated field until
// a boundary is reached).  The reference takes a user callback sub_int(ndim, nvar, field, y, dy_var)
// (:61-74); a batched device version cannot call back into Fortran, so the integrand is one of
//   0: none (nvar = 0)
//   1: dy_var(:) = 1            -- arc length, exactly test_trace_field.f90:57-64
//   2: dy_var(:) = norm2(field) -- potential-like line integral of |field|
// y is (ndim+nvar, max_steps), y_field is (ndim, max_steps), both column-major as in Fortran.
// i_icell_mask = 0 <=> both optional mask arguments absent.  boundary_material may be null.
void sub_int_builtin(int integrand, int ndim, int nvar, const double* field, double* dy_var) {
  double v = 1.0;
  if (integrand == 2) v = norm2(V3{field[0], field[1], ndim > 2 ? field[2] : 0.0});
  for (int i = 0; i < nvar; ++i) dy_var[i] = v;
}

// smallest |err - 1| over all accept / reject decisions (:1165) of the latest batch: how far the closest decision was from
// flipping.  The CUDA tracer's curves agree to 1e-10 (pow() differs in the last bit), so a margin far above that means
// no decision of the batch can have gone the other way on the device.
double g_trace_margin = 1e300;

void integrate_along_field(const Grid& g, int ndim, int nvar, int integrand, const int* i_field, double min_dx,
                           double max_dx, int max_steps, double rtol, double atol, bool reverse, double* y,
                           double* y_field, int& n_steps, bool axisymmetric, int i_icell_mask, int mask_value,
                           int* boundary_material, int nn_mode) {
  const double safety_fac = 0.8, inv_24 = 1 / 24.0, inv_9 = 1 / 9.0, min_radius = 1e-12, eps = 1e-8;
  double line_margin = 1e300;
  struct MarginGuard {
    double& m;
    ~MarginGuard() {
#pragma omp critical(iu_trace_margin)
      if (m < g_trace_margin) g_trace_margin = m;
    }
  } margin_guard{line_margin};
  const int nt = ndim + nvar;
  constexpr int MAXT = 8;
  double ys[MAXT], y_2nd[MAXT], k[4][MAXT], field[3], scales[MAXT];
  double r0[3] = {0, 0, 0}, r[3] = {0, 0, 0};
  V3 r_p{0, 0, 0};
  auto Y = [&](int i, int step) -> double& { return y[(size_t)(step - 1) * nt + i]; };         // y(i+1, step)
  auto YF = [&](int i, int step) -> double& { return y_field[(size_t)(step - 1) * ndim + i]; };
  auto mask_of = [&](int ic) { return g.icell_data[(size_t)(i_icell_mask - 1) * g.n_cells + (ic - 1)]; };
  auto set_boundary = [&](int ic) {
    if (boundary_material) *boundary_material = ic <= 0 ? -1 : mask_of(ic);                      // :1062-1068, :1087-1093
  };
  auto unitvec = [&](const double* f, double* out) {                                             // get_unitvec, :1192-1201
    const double nrm = norm2(V3{f[0], f[1], ndim > 2 ? f[2] : 0.0});
    for (int i = 0; i < ndim; ++i) { out[i] = f[i] / nrm; if (reverse) out[i] = -out[i]; }
  };
  n_steps = 1;
  for (int i = 0; i < ndim; ++i) r0[i] = Y(i, 1);
  int last_rejected = -100;
  double dx = max_dx;
  int i_cell = 0;
  // field at the initial point: iu_interpolate_at, cold (:1054)
  get_cell(g, r0, i_cell, nn_mode, nullptr, nullptr);
  if (i_cell > 0) interpolate_at_icell(g, r0, ndim, i_field, field, i_cell);
  bool valid = i_cell > 0 && (i_icell_mask <= 0 || mask_of(i_cell) == mask_value);             // cell_is_valid, :1203-1215
  if (!valid) { set_boundary(i_cell); return; }
  for (int i = 0; i < ndim; ++i) YF(i, 1) = field[i];
  int i_cell_prev = i_cell, status = 0, n_its = 0;
  for (long iteration = 1;; ++iteration) {
    if (status != 0) {                                                                           // :1081-1097
      last_rejected = (int)std::min<long>(iteration - 1, 2000000000L);
      dx = (1 - eps) * norm2(V3{r_p.x - r0[0], r_p.y - r0[1], r_p.z - r0[2]});
      if (dx < min_dx) { set_boundary(i_cell); return; }
    }
    i_cell = i_cell_prev;
    for (int i = 0; i < ndim; ++i) r0[i] = Y(i, n_steps);
    for (int i = 0; i < nt; ++i) ys[i] = Y(i, n_steps);
    for (int i = 0; i < ndim; ++i) field[i] = YF(i, n_steps);
    unitvec(field, k[0]);
    sub_int_builtin(integrand, ndim, nvar, field, k[0] + ndim);
    auto substep = [&](int stage) -> bool {      // position ys -> r, walk, interpolate; false => cycle
      for (int i = 0; i < ndim; ++i) r[i] = ys[i];
      if (axisymmetric) r[0] = std::max(r[0], min_radius);
      walk(g, std::numeric_limits<int>::max() - 1, {r0[0], r0[1], r0[2]}, {r[0], r[1], r[2]}, i_cell_prev, i_cell, r_p,
           n_its, status, i_icell_mask, nullptr);
      if (status != 0) return false;
      interpolate_at_icell(g, r, ndim, i_field, field, i_cell);
      unitvec(field, k[stage]);
      sub_int_builtin(integrand, ndim, nvar, field, k[stage] + ndim);
      return true;
    };
    { const double f = 0.5 * dx; for (int i = 0; i < nt; ++i) ys[i] = Y(i, n_steps) + f * k[0][i]; }      // :1118
    if (!substep(1)) continue;
    { const double f = 0.75 * dx; for (int i = 0; i < nt; ++i) ys[i] = Y(i, n_steps) + f * k[1][i]; }     // :1131
    if (!substep(2)) continue;
    { const double f = dx * inv_9;                                                                        // :1144-1145
      for (int i = 0; i < nt; ++i) ys[i] = Y(i, n_steps) + f * ((2 * k[0][i] + 3 * k[1][i]) + 4 * k[2][i]); }
    if (!substep(3)) continue;
    { const double f = dx * inv_24;                                                                       // :1159-1160
      for (int i = 0; i < nt; ++i) y_2nd[i] = Y(i, n_steps) + f * (((7 * k[0][i] + 6 * k[1][i]) + 8 * k[2][i]) + 3 * k[3][i]); }
    double sum = 0;
    for (int i = 0; i < nt; ++i) {
      scales[i] = atol + std::max(std::fabs(ys[i]), std::fabs(y_2nd[i])) * rtol;
      const double e = (ys[i] - y_2nd[i]) / scales[i];
      sum += e * e;
    }
    const double err = std::sqrt(sum / 3);
    if (!(dx < 2 * min_dx) && std::fabs(err - 1) < line_margin) line_margin = std::fabs(err - 1);
    if (err <= 1 || dx < 2 * min_dx) {                                                                    // :1165-1176
      n_steps = n_steps + 1;
      if (n_steps > max_steps) return;
      for (int i = 0; i < nt; ++i) Y(i, n_steps) = ys[i];
      if (axisymmetric) Y(0, n_steps) = std::max(Y(0, n_steps), min_radius);
      for (int i = 0; i < ndim; ++i) YF(i, n_steps) = field[i];
      i_cell_prev = i_cell;
    } else {
      last_rejected = (int)std::min<long>(iteration, 2000000000L);
    }
    const double max_growth = last_rejected > iteration - 2 ? 1.0 : 2.0;                                  // :1178-1184
    const double dx_factor = std::min(max_growth, safety_fac * std::pow(1 / err, 1 / 3.0));
    dx = std::max(min_dx, std::min(max_dx, dx * dx_factor));
  }
}

// ---------------------------------------------------------------------------------
// BINDA reader, m_binda.f90:52-194 (format written by convert_to_binary.py:68-115)
// ---------------------------------------------------------------------------------
struct BindaEntry { std::string name, dtype, meta; int64_t ndim; int64_t shape[8]; int64_t offset; };
std::string trim128(const char* s) {
  int n = 128;
  while (n > 0 && (s[n - 1] == ' ' || s[n - 1] == '\0')) --n;
  return std::string(s, n);
}
bool binda_read(const std::string& path, std::vector<BindaEntry>& ent, std::vector<char>& buf) {
  std::ifstream f(path, std::ios::binary);
  if (!f) { g_err = "cannot open " + path; return false; }
  buf.assign(std::istreambuf_iterator<char>(f), std::istreambuf_iterator<char>());
  if (buf.size() < 24 || std::string(buf.data(), 5) != "BINDA") { g_err = "Wrong file format"; return false; }
  int64_t n_entries; std::memcpy(&n_entries, &buf[8], 8);
  size_t pos = 24;
  for (int64_t i = 0; i < n_entries; ++i) {
    BindaEntry e;
    e.name = trim128(&buf[pos]); e.dtype = trim128(&buf[pos + 128]); e.meta = trim128(&buf[pos + 256]);
    std::memcpy(&e.ndim, &buf[pos + 384], 8);
    std::memcpy(e.shape, &buf[pos + 392], 64);
    std::memcpy(&e.offset, &buf[pos + 456], 8);
    pos += 464;
    ent.push_back(e);
  }
  return true;
}
template <class T> bool read_int(const BindaEntry& e, const std::vector<char>& buf, size_t n, T* out) {
  if (e.dtype == "int32") { const int32_t* s = (const int32_t*)&buf[e.offset]; for (size_t i = 0; i < n; ++i) out[i] = (T)s[i]; }
  else if (e.dtype == "int64") { const int64_t* s = (const int64_t*)&buf[e.offset]; for (size_t i = 0; i < n; ++i) out[i] = (T)s[i]; }
  else { g_err = "Unsupported data type " + e.dtype; return false; }
  return true;
}
bool read_f64(const BindaEntry& e, const std::vector<char>& buf, size_t n, double* out) {
  if (e.dtype == "float64") std::memcpy(out, &buf[e.offset], n * 8);
  else if (e.dtype == "float32") { const float* s = (const float*)&buf[e.offset]; for (size_t i = 0; i < n; ++i) out[i] = (double)s[i]; }
  else { g_err = "Unsupported data type " + e.dtype; return false; }
  return true;
}

}  // namespace

// =====================================================================================
// C interface (ctypes)
// =====================================================================================
extern "C" {

const char* iu_oracle_last_error() { return g_err.c_str(); }

// cells / neighbors 1-based (as after iu_read_grid's +1, :871-872); point_data (Np, n_pd)
// column-major.
void* iu_oracle_grid_create(int cell_type, int n_points, int n_cells, const double* points, const int* cells,
                            const int* neighbors, int n_pd, const double* point_data, int n_cd,
                            const double* cell_data, int n_icd, const int* icell_data) {
  auto g = std::make_unique<Grid>();
  g->cell_type = cell_type;
  g->ppc = cell_type == IU_TRIANGLE ? 3 : 4;
  g->n_points = n_points; g->n_cells = n_cells;
  g->points.assign(points, points + (size_t)3 * n_points);
  g->cells.assign(cells, cells + (size_t)g->ppc * n_cells);
  g->neighbors.assign(neighbors, neighbors + (size_t)g->ppc * n_cells);
  g->n_point_data = n_pd; g->n_cell_data = n_cd; g->n_icell_data = n_icd;
  if (n_pd) g->point_data.assign(point_data, point_data + (size_t)n_pd * n_points);
  if (n_cd) g->cell_data.assign(cell_data, cell_data + (size_t)n_cd * n_cells);
  if (n_icd) g->icell_data.assign(icell_data, icell_data + (size_t)n_icd * n_cells);
  finish_grid(*g);
  return g.release();
}

// iu_read_grid, :820-927 (binda input only; the .vtu -> .binda subprocess is out of scope)
void* iu_oracle_grid_from_binda(const char* path, double coord_scale_factor) {
  std::vector<BindaEntry> ent; std::vector<char> buf;
  if (!binda_read(path, ent, buf)) return nullptr;
  auto find = [&](const char* nm) { for (size_t i = 0; i < ent.size(); ++i) if (ent[i].name == nm) return (int)i; return -1; };
  auto g = std::make_unique<Grid>();
  int ix = find("cells");
  if (ix < 0) { g_err = "cells not found in binda file"; return nullptr; }
  if (ent[ix].ndim != 2) { g_err = "Invalid ndim"; return nullptr; }
  g->n_cells = (int)ent[ix].shape[0]; g->ppc = (int)ent[ix].shape[1];
  g->cells.resize((size_t)g->ppc * g->n_cells);
  if (!read_int(ent[ix], buf, g->cells.size(), g->cells.data())) return nullptr;
  if (ent[ix].meta == "triangle") g->cell_type = IU_TRIANGLE;
  else if (ent[ix].meta == "quad") g->cell_type = IU_QUAD;
  else if (ent[ix].meta == "tetra") g->cell_type = IU_TETRA;
  else { g_err = "Unsupported cell type"; return nullptr; }
  ix = find("points");
  if (ix < 0) { g_err = "points not found in binda file"; return nullptr; }
  g->n_points = (int)ent[ix].shape[0];
  g->points.resize((size_t)3 * g->n_points);
  if (!read_f64(ent[ix], buf, g->points.size(), g->points.data())) return nullptr;
  ix = find("cell_neighbors");
  if (ix < 0) { g_err = "cell_neighbors not found in binda file"; return nullptr; }
  g->neighbors.resize(g->cells.size());
  if (!read_int(ent[ix], buf, g->neighbors.size(), g->neighbors.data())) return nullptr;
  for (auto& x : g->points) x = x * coord_scale_factor;   // :858-860 (factor 1 is exact)
  for (auto& c : g->cells) c += 1;                         // :871
  for (auto& c : g->neighbors) c += 1;                     // :872
  for (auto& e : ent) {
    if (e.name == "point_data") {
      g->point_names.push_back(e.meta);
      g->point_data.resize((size_t)(++g->n_point_data) * g->n_points);
      if (!read_f64(e, buf, g->n_points, &g->point_data[(size_t)(g->n_point_data - 1) * g->n_points])) return nullptr;
    } else if (e.name == "cell_data") {
      g->cell_names.push_back(e.meta);
      g->cell_data.resize((size_t)(++g->n_cell_data) * g->n_cells);
      if (!read_f64(e, buf, g->n_cells, &g->cell_data[(size_t)(g->n_cell_data - 1) * g->n_cells])) return nullptr;
    } else if (e.name == "icell_data") {
      g->icell_names.push_back(e.meta);
      g->icell_data.resize((size_t)(++g->n_icell_data) * g->n_cells);
      if (!read_int(e, buf, g->n_cells, &g->icell_data[(size_t)(g->n_icell_data - 1) * g->n_cells])) return nullptr;
    }
  }
  finish_grid(*g);
  return g.release();
}

void iu_oracle_grid_destroy(void* h) { delete (Grid*)h; }

// info[0..6] = cell_type, n_points, n_cells, ppc, n_point_data, n_cell_data, n_icell_data
void iu_oracle_grid_info(void* h, int* info) {
  Grid& g = *(Grid*)h;
  info[0] = g.cell_type; info[1] = g.n_points; info[2] = g.n_cells; info[3] = g.ppc;
  info[4] = g.n_point_data; info[5] = g.n_cell_data; info[6] = g.n_icell_data;
}
// which: 0 points, 1 cell_points, 2 normals, 3 volume, 4 centers, 5 point_data, 6 cell_data
const double* iu_oracle_f64_array(void* h, int which, int64_t* n) {
  Grid& g = *(Grid*)h;
  const std::vector<double>* v = nullptr;
  switch (which) {
    case 0: v = &g.points; break; case 1: v = &g.cell_points; break; case 2: v = &g.normals; break;
    case 3: v = &g.cell_volume; break; case 4: v = &g.centers; break; case 5: v = &g.point_data; break;
    case 6: v = &g.cell_data; break; default: *n = 0; return nullptr;
  }
  *n = (int64_t)v->size(); return v->data();
}
// which: 0 cells, 1 neighbors, 2 icell_data
const int* iu_oracle_i32_array(void* h, int which, int64_t* n) {
  Grid& g = *(Grid*)h;
  const std::vector<int>* v = which == 0 ? &g.cells : which == 1 ? &g.neighbors : &g.icell_data;
  *n = (int64_t)v->size(); return v->data();
}
// iu_get_point_data_index etc., :106-116, :224-247.  kind 0 point, 1 cell, 2 icell. -1 if absent.
int iu_oracle_data_index(void* h, int kind, const char* name) {
  Grid& g = *(Grid*)h;
  const auto& v = kind == 0 ? g.point_names : kind == 1 ? g.cell_names : g.icell_names;
  for (size_t i = 0; i < v.size(); ++i) if (v[i] == name) return (int)i + 1;
  return -1;
}
// overwrite a point_data column (callers mutate ug%point_data: test_tetra.f90:38-40)
void iu_oracle_set_point_data(void* h, int col, const double* data) {
  Grid& g = *(Grid*)h;
  if ((int)(g.point_data.size() / std::max(1, g.n_points)) < col) { g.point_data.resize((size_t)col * g.n_points); g.n_point_data = col; }
  std::memcpy(&g.point_data[(size_t)(col - 1) * g.n_points], data, (size_t)g.n_points * 8);
}

static int eff_threads(int nthreads) {
#ifdef _OPENMP
  return nthreads > 0 ? nthreads : omp_get_max_threads();
#else
  (void)nthreads; return 1;
#endif
}

// nearest centre for n queries.  mode 0 kdtree2 restatement, 1 brute force (also fills
// best/second squared distances when non-null).  idx 1-based.
void iu_oracle_nearest(void* h, int64_t n, const double* r, int mode, int* idx, double* best, double* second, int nthreads) {
  const Grid& g = *(Grid*)h;
  int nt = eff_threads(nthreads); (void)nt;
#pragma omp parallel for num_threads(nt) schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    if (mode == 0) idx[i] = find_nearby_cell_kdtree(g, r + 3 * i);
    else idx[i] = find_nearby_cell_brute(g, r + 3 * i, best ? best + i : nullptr, second ? second + i : nullptr);
  }
}

// batched iu_get_cell, :414-434.  n_steps / margins optional (may be null).
// margins[2*i] = min |distance_left| over decisions (:704), margins[2*i+1] = min gap between
// best and second-best face distance (:755); +inf when not applicable.
void iu_oracle_get_cell(void* h, int64_t n, const double* r, int* i_cell, int nn_mode, int* n_steps, double* margins, int nthreads) {
  const Grid& g = *(Grid*)h;
  int nt = eff_threads(nthreads); (void)nt;
#pragma omp parallel for num_threads(nt) schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    WalkDiag d{std::numeric_limits<double>::infinity(), std::numeric_limits<double>::infinity()};
    int ns = 0;
    get_cell(g, r + 3 * i, i_cell[i], nn_mode, &ns, margins ? &d : nullptr);
    if (n_steps) n_steps[i] = ns;
    if (margins) { margins[2 * i] = d.min_abs_left; margins[2 * i + 1] = d.min_face_gap; }
  }
}

// batched iu_interpolate_at, :480-495.  var_at_r is (n_vars, n) column-major; rows of
// not-found queries are left untouched.  Returns -1 (and does nothing) if any hint > n_cells (:490).
int iu_oracle_interpolate_at(void* h, int64_t n, const double* r, int n_vars, const int* i_vars, double* var_at_r,
                             int* i_cell, int nn_mode, int nthreads) {
  const Grid& g = *(Grid*)h;
  for (int64_t i = 0; i < n; ++i) if (i_cell[i] > g.n_cells) { g_err = "i_cell > ug%n_cells"; return -1; }
  int nt = eff_threads(nthreads); (void)nt;
#pragma omp parallel for num_threads(nt) schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    get_cell(g, r + 3 * i, i_cell[i], nn_mode, nullptr, nullptr);
    if (i_cell[i] <= 0) continue;
    interpolate_at_icell(g, r + 3 * i, n_vars, i_vars, var_at_r + (size_t)n_vars * i, i_cell[i]);
  }
  return 0;
}

// batched iu_interpolate_at_icell, :498-527 (cells given, all must be valid)
void iu_oracle_interpolate_at_icell(void* h, int64_t n, const double* r, int n_vars, const int* i_vars, double* var_at_r,
                                    const int* i_cell) {
  const Grid& g = *(Grid*)h;
  for (int64_t i = 0; i < n; ++i)
    if (i_cell[i] >= 1) interpolate_at_icell(g, r + 3 * i, n_vars, i_vars, var_at_r + (size_t)n_vars * i, i_cell[i]);
}

// batched iu_get_cell_scalar_at / iu_get_icell_scalar_at, :437-462
void iu_oracle_get_cell_scalar_at(void* h, int64_t n, const double* r, int i_var, int* i_guess, double* res, int nn_mode) {
  const Grid& g = *(Grid*)h;
  for (int64_t i = 0; i < n; ++i) {
    get_cell(g, r + 3 * i, i_guess[i], nn_mode, nullptr, nullptr);
    if (i_guess[i] >= 1) res[i] = g.cell_data[(size_t)(i_var - 1) * g.n_cells + (i_guess[i] - 1)];
  }
}
// the same over OpenMP threads (bench.py's CPU leg of config C5)
void iu_oracle_get_cell_scalar_at_mt(void* h, int64_t n, const double* r, int i_var, int* i_guess, double* res, int nn_mode,
                                     int nthreads) {
  const Grid& g = *(Grid*)h;
  int nt = eff_threads(nthreads); (void)nt;
#pragma omp parallel for num_threads(nt) schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    get_cell(g, r + 3 * i, i_guess[i], nn_mode, nullptr, nullptr);
    if (i_guess[i] >= 1) res[i] = g.cell_data[(size_t)(i_var - 1) * g.n_cells + (i_guess[i] - 1)];
  }
}
void iu_oracle_get_icell_scalar_at(void* h, int64_t n, const double* r, int i_var, int* i_guess, int* res, int nn_mode) {
  const Grid& g = *(Grid*)h;
  for (int64_t i = 0; i < n; ++i) {
    get_cell(g, r + 3 * i, i_guess[i], nn_mode, nullptr, nullptr);
    if (i_guess[i] >= 1) res[i] = g.icell_data[(size_t)(i_var - 1) * g.n_cells + (i_guess[i] - 1)];
  }
}

// batched iu_get_cell_through_neighbors, :667-725.  i_icell_mask = 0 => absent.
// On the early return (|r1-r0| < 1e-100) r_p and n_steps are left untouched, as in the reference.
void iu_oracle_walk(void* h, int64_t n, int max_steps, const double* r0, const double* r1, const int* ic0, int* ic1,
                    double* r_p, int* n_steps, int* status, int i_icell_mask) {
  const Grid& g = *(Grid*)h;
  for (int64_t i = 0; i < n; ++i) {
    V3 rp{r_p[3 * i], r_p[3 * i + 1], r_p[3 * i + 2]};
    int ns = n_steps[i];
    walk(g, max_steps, {r0[3 * i], r0[3 * i + 1], r0[3 * i + 2]}, {r1[3 * i], r1[3 * i + 1], r1[3 * i + 2]}, ic0[i],
         ic1[i], rp, ns, status[i], i_icell_mask, nullptr);
    r_p[3 * i] = rp.x; r_p[3 * i + 1] = rp.y; r_p[3 * i + 2] = rp.z;
    n_steps[i] = ns;
  }
}

void iu_oracle_point_is_inside_cell(void* h, int64_t n, const double* r, const int* i_cell, int* inside) {
  const Grid& g = *(Grid*)h;
  for (int64_t i = 0; i < n; ++i) inside[i] = (i_cell[i] >= 1 && point_is_inside_cell(g, r + 3 * i, i_cell[i])) ? 1 : 0;
}

// batched iu_integrate_along_field (built-in integrands, see integrate_along_field above).
// y: (ndim+nvar, max_steps, n) inout, y_field: (ndim, max_steps, n) inout, n_steps[n] out,
// boundary_material[n] inout or null.
void iu_oracle_integrate_along_field(void* h, int64_t n, int ndim, int nvar, int integrand, const int* i_field, double min_dx,
                                     double max_dx, int max_steps, double rtol, double atol, int reverse, double* y,
                                     double* y_field, int* n_steps, int axisymmetric, int i_icell_mask, int mask_value,
                                     int* boundary_material, int nn_mode, int nthreads) {
  const Grid& g = *(Grid*)h;
  const size_t ny = (size_t)(ndim + nvar) * max_steps, nf = (size_t)ndim * max_steps;
  int nt = eff_threads(nthreads); (void)nt;
  g_trace_margin = 1e300;
#pragma omp parallel for num_threads(nt) schedule(dynamic, 64)
  for (int64_t i = 0; i < n; ++i)
    integrate_along_field(g, ndim, nvar, integrand, i_field, min_dx, max_dx, max_steps, rtol, atol, reverse != 0, y + ny * i,
                          y_field + nf * i, n_steps[i], axisymmetric != 0, i_icell_mask, mask_value,
                          boundary_material ? boundary_material + i : nullptr, nn_mode);
}

// get_cell_neighbors, convert_to_binary.py:118-162.  cells0 (n_cells, ppc) 0-based as on disk; neigh0 (n_cells, ppc)
// receives the 0-based neighbour behind face i = vertices cell[(i + k) % ppc], k = 0..n_points_face-1, or -1.
// Coincident points are merged first (np.unique(points, axis=0), :130: only the equivalence classes matter, so the
// class id here is the lowest point index of the class, not numpy's rank).  The reference's dict of sorted face
// tuples (:139-160) is restated as a sort of (face, cell) records: a face with exactly two cells links them, any
// other count leaves -1, which is what `len(all_cells) == 2` does.
int iu_oracle_cell_neighbors(int n_points, int n_cells, int ppc, int n_points_face, const double* points, const int* cells0,
                             int* neigh0) {
  if (ppc < 3 || ppc > 4 || n_points_face < 2 || n_points_face > 3) { g_err = "unsupported cell shape"; return -1; }
  std::vector<int> order(n_points), cls(n_points);
  for (int i = 0; i < n_points; ++i) order[i] = i;
  auto key_less = [&](int a, int b) {
    const double* pa = points + 3 * (size_t)a; const double* pb = points + 3 * (size_t)b;
    for (int k = 0; k < 3; ++k) { if (pa[k] < pb[k]) return true; if (pa[k] > pb[k]) return false; }
    return a < b;
  };
  std::sort(order.begin(), order.end(), key_less);
  for (int j = 0; j < n_points; ++j) {
    const int i = order[j];
    bool same = false;
    if (j > 0) {
      const double* pa = points + 3 * (size_t)order[j - 1]; const double* pb = points + 3 * (size_t)i;
      same = pa[0] == pb[0] && pa[1] == pb[1] && pa[2] == pb[2];
    }
    cls[i] = same ? cls[order[j - 1]] : i;
  }
  struct Face { uint32_t v[3]; uint32_t slot; };   // slot = cell * ppc + face index
  const size_t nf = (size_t)n_cells * ppc;
  std::vector<Face> faces(nf);
  for (size_t c = 0; c < (size_t)n_cells; ++c)
    for (int i = 0; i < ppc; ++i) {
      Face f; f.v[2] = 0; f.slot = (uint32_t)(c * ppc + i);
      for (int k = 0; k < n_points_face; ++k) f.v[k] = (uint32_t)cls[cells0[c * ppc + (i + k) % ppc]];
      std::sort(f.v, f.v + n_points_face);
      faces[c * ppc + i] = f;
    }
  std::sort(faces.begin(), faces.end(), [](const Face& a, const Face& b) {
    for (int k = 0; k < 3; ++k) if (a.v[k] != b.v[k]) return a.v[k] < b.v[k];
    return a.slot < b.slot;
  });
  for (size_t i = 0; i < nf; ++i) neigh0[i] = -1;
  for (size_t i = 0; i < nf;) {
    size_t j = i + 1;
    while (j < nf && faces[j].v[0] == faces[i].v[0] && faces[j].v[1] == faces[i].v[1] && faces[j].v[2] == faces[i].v[2]) ++j;
    if (j - i == 2) {
      neigh0[faces[i].slot] = (int)(faces[i + 1].slot / ppc);
      neigh0[faces[i + 1].slot] = (int)(faces[i].slot / ppc);
    }
    i = j;
  }
  return 0;
}

// smallest |err - 1| over the accept / reject decisions of the latest iu_oracle_integrate_along_field batch
double iu_oracle_last_trace_margin() { return g_trace_margin; }

int iu_oracle_max_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

// benchmark.f90:46-55 / :60-69: one timed pass of iu_interpolate_at over n queries
// (i_cell in/out: 0 = cold pass, previous cells = warm pass).  Returns seconds.
double iu_oracle_benchmark_pass(void* h, int64_t n, const double* r, int n_vars, const int* i_vars, double* res,
                                int* i_cell, int nthreads) {
  double t0, t1;
#ifdef _OPENMP
  t0 = omp_get_wtime();
#else
  t0 = (double)clock() / CLOCKS_PER_SEC;
#endif
  iu_oracle_interpolate_at(h, n, r, n_vars, i_vars, res, i_cell, 0, nthreads);
#ifdef _OPENMP
  t1 = omp_get_wtime();
#else
  t1 = (double)clock() / CLOCKS_PER_SEC;
#endif
  return t1 - t0;
}

}  // extern "C"
