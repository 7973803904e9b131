// Throw-away design experiment (not product, not oracle): count node/leaf/point visits of
// exact-NN traversals on a balanced implicit kd-tree over Kuhn-tet centres.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <random>
#include <vector>
struct P { double c[3]; int id; };
static std::vector<P> pts;
static std::vector<double> split; static std::vector<uint8_t> sdim; static std::vector<int> leaf_start;
static int D;
static void build(int node, int level, int lo, int hi) {
  if (level == D) { leaf_start[node - ((1 << D) - 1)] = lo; return; }
  double mn[3] = {1e300, 1e300, 1e300}, mx[3] = {-1e300, -1e300, -1e300};
  for (int i = lo; i < hi; ++i) for (int k = 0; k < 3; ++k) { mn[k] = std::min(mn[k], pts[i].c[k]); mx[k] = std::max(mx[k], pts[i].c[k]); }
  int d = 0; for (int k = 1; k < 3; ++k) if (mx[k] - mn[k] > mx[d] - mn[d]) d = k;
  int mid = lo + (hi - lo + 1) / 2;
  std::nth_element(pts.begin() + lo, pts.begin() + mid, pts.begin() + hi, [d](const P& a, const P& b) { return a.c[d] < b.c[d]; });
  split[node] = pts[mid].c[d]; sdim[node] = d;
  build(2 * node + 1, level + 1, lo, mid); build(2 * node + 2, level + 1, mid, hi);
}
static long n_nodes, n_leaves, n_points;
static double best; static int besti; static const double* Q;
static void scan(int leaf) {
  ++n_leaves;
  for (int i = leaf_start[leaf]; i < leaf_start[leaf + 1]; ++i) {
    ++n_points;
    double dx = pts[i].c[0] - Q[0], dy = pts[i].c[1] - Q[1], dz = pts[i].c[2] - Q[2];
    double sd = dx * dx + dy * dy + dz * dz;
    if (sd < best) { best = sd; besti = i; }
  }
}
static void plane_only(int node, int level) {
  if (level == D) { scan(node - ((1 << D) - 1)); return; }
  ++n_nodes;
  double d = Q[sdim[node]] - split[node];
  int near = d < 0 ? 2 * node + 1 : 2 * node + 2, far = d < 0 ? 2 * node + 2 : 2 * node + 1;
  plane_only(near, level + 1);
  if (d * d <= best) plane_only(far, level + 1);
}
static void incr(int node, int level, double ox, double oy, double oz) {
  if (level == D) { scan(node - ((1 << D) - 1)); return; }
  ++n_nodes;
  int k = sdim[node];
  double d = Q[k] - split[node];
  int near = d < 0 ? 2 * node + 1 : 2 * node + 2, far = d < 0 ? 2 * node + 2 : 2 * node + 1;
  incr(near, level + 1, ox, oy, oz);
  double o[3] = {ox, oy, oz}; o[k] = d;
  double rd = o[0] * o[0] + o[1] * o[1] + o[2] * o[2];
  if (rd <= best) incr(far, level + 1, o[0], o[1], o[2]);
}
int main(int argc, char** argv) {
  int n = argc > 1 ? atoi(argv[1]) : 48; int leafmax = argc > 2 ? atoi(argv[2]) : 8;
  double h = 1.0 / n; std::mt19937_64 rng(1); std::uniform_real_distribution<double> U(-0.1 * h, 0.1 * h), U01(0.1, 0.9);
  std::vector<double> v((size_t)(n + 1) * (n + 1) * (n + 1) * 3);
  auto vid = [&](int i, int j, int k) { return ((size_t)k * (n + 1) + j) * (n + 1) + i; };
  for (int k = 0; k <= n; ++k) for (int j = 0; j <= n; ++j) for (int i = 0; i <= n; ++i) {
    bool interior = i > 0 && i < n && j > 0 && j < n && k > 0 && k < n; size_t a = vid(i, j, k) * 3;
    v[a] = i * h + (interior ? U(rng) : 0); v[a + 1] = j * h + (interior ? U(rng) : 0); v[a + 2] = k * h + (interior ? U(rng) : 0);
  }
  int perms[6][3] = {{0,1,2},{0,2,1},{1,0,2},{1,2,0},{2,0,1},{2,1,0}};
  for (int k = 0; k < n; ++k) for (int j = 0; j < n; ++j) for (int i = 0; i < n; ++i) for (auto& pm : perms) {
    int a[3] = {i, j, k}; double c[3] = {0, 0, 0};
    for (int s = 0; s < 4; ++s) { size_t id = vid(a[0], a[1], a[2]) * 3; for (int t = 0; t < 3; ++t) c[t] += v[id + t]; if (s < 3) a[pm[s]]++; }
    P p; for (int t = 0; t < 3; ++t) p.c[t] = c[t] / 4; p.id = (int)pts.size(); pts.push_back(p);
  }
  int N = (int)pts.size(); D = 0; while ((N + (1 << D) - 1) / (1 << D) > leafmax) ++D;
  split.resize((1 << D) - 1 + 1); sdim.resize((1 << D)); leaf_start.assign((1 << D) + 1, N);
  build(0, 0, 0, N);
  printf("N=%d D=%d leaves=%d avg/leaf=%.2f\n", N, D, 1 << D, (double)N / (1 << D));
  int nq = 200000; std::vector<double> q(3 * nq); for (auto& x : q) x = U01(rng);
  for (int mode = 0; mode < 2; ++mode) {
    n_nodes = n_leaves = n_points = 0; long chk = 0;
    for (int i = 0; i < nq; ++i) { Q = &q[3 * i]; best = 1e300; besti = -1; if (mode == 0) plane_only(0, 0); else incr(0, 0, 0, 0, 0); chk += besti; }
    printf("%s: nodes/q=%.1f leaves/q=%.2f points/q=%.1f chk=%ld\n", mode ? "incremental" : "plane-only", (double)n_nodes / nq, (double)n_leaves / nq, (double)n_points / nq, chk);
  }
}
