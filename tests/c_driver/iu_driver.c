/* Plain-C driver of include/iu_b200.h: the reference's own test programs restated against the C ABI, the way a
 * non-Fortran, non-Python host would call it.
 *
 *   iu_driver <dir with triangle.binda quad.binda tetra.binda>
 *
 * test_triangle.f90 / test_quad.f90 / test_tetra.f90: 1000 random points in the bounding box, the point variable
 * 'Polynomial' (x + y (+ z) + 1; test_tetra.f90:37-40 overwrites it with the analytic solution first), every point
 * queried on its own with i_cell = 0 like the reference's loop (:46-49), |res - solution| <= 1e-14 (:23).
 * Then benchmark.f90's flow on the triangle grid through the batch call: cold pass, warm pass with the found cells.
 * Exit status 0 and "PASSED" on success, like the reference programs.  Built by __graft_entry__.build():
 *   gcc -std=c99 -O2 -I include tests/c_driver/iu_driver.c -L interpolate_unstructured_b200/lib -liu_b200 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "iu_b200.h"

#define CHECK(call)                                                                    \
  do {                                                                                 \
    int rc_ = (call);                                                                  \
    if (rc_ != IU_B200_OK) {                                                           \
      fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, iu_b200_last_error());       \
      exit(1);                                                                         \
    }                                                                                  \
  } while (0)

static uint64_t rng_state = 0x243F6A8885A308D3ULL;
static double uniform01(void) { /* xorshift64*: any generator will do (the reference uses random_number) */
  rng_state ^= rng_state >> 12; rng_state ^= rng_state << 25; rng_state ^= rng_state >> 27;
  return (double)((rng_state * 0x2545F4914F6CDD1DULL) >> 11) * (1.0 / 9007199254740992.0);
}

static int test_interpolation(const char* dir, const char* name, int ndim, int overwrite_solution) {
  char path[4096];
  iu_b200_grid* ug = NULL;
  int32_t info[8];
  double rmin[3], rmax[3];
  const int n_samples = 1000;
  const double threshold = 1e-14;
  snprintf(path, sizeof path, "%s/%s.binda", dir, name);
  CHECK(iu_b200_grid_from_binda(path, 1.0, -1, &ug));
  CHECK(iu_b200_grid_info(ug, info));
  CHECK(iu_b200_grid_bounds(ug, rmin, rmax));
  const int ivar = iu_b200_get_point_data_index(ug, "Polynomial");
  if (ivar == -1) { fprintf(stderr, "Point data 'Polynomial' not found\n"); return 1; }
  if (overwrite_solution) { /* test_tetra.f90:37-40: ug%point_data(n, ivar) = solution(ug%points(:, n)) */
    const int np = info[1];
    double* pts = (double*)malloc(sizeof(double) * 3 * (size_t)np);
    double* col = (double*)malloc(sizeof(double) * (size_t)np);
    CHECK(iu_b200_export_mesh(ug, pts, NULL));
    for (int n = 0; n < np; ++n) col[n] = pts[3 * n] + pts[3 * n + 1] + pts[3 * n + 2] + 1;
    CHECK(iu_b200_update_point_data(ug, ivar, 1, col));
    free(pts); free(col);
  }
  for (int n = 0; n < n_samples; ++n) {
    double r[3], res = -1e300;
    int32_t i_cell = 0;
    for (int k = 0; k < 3; ++k) r[k] = rmin[k] + uniform01() * (rmax[k] - rmin[k]);
    if (ndim == 2) r[2] = rmin[2];
    CHECK(iu_b200_interpolate_scalar_at_batch(ug, 1, r, ivar, &res, &i_cell));   /* iu_interpolate_scalar_at, one point */
    if (i_cell < 1) { fprintf(stderr, "%s: point %d not located\n", name, n); return 1; }
    const double difference = fabs(res - (r[0] + r[1] + r[2] + 1));
    if (difference > threshold) {
      fprintf(stderr, "%s: Difference of %g at %g %g %g\nDifference exceeds threshold\n", name, difference, r[0], r[1], r[2]);
      return 1;
    }
  }
  CHECK(iu_b200_grid_destroy(ug));
  printf(" %s PASSED\n", name);
  return 0;
}

static int benchmark_flow(const char* dir) { /* benchmark.f90:14-71 with the batch call */
  char path[4096];
  iu_b200_grid* ug = NULL;
  double rmin[3], rmax[3];
  const int64_t n = 100000;
  snprintf(path, sizeof path, "%s/triangle.binda", dir);
  CHECK(iu_b200_grid_from_binda(path, 1.0, -1, &ug));
  CHECK(iu_b200_grid_bounds(ug, rmin, rmax));
  double* r = (double*)malloc(sizeof(double) * 3 * (size_t)n);
  double* res = (double*)malloc(sizeof(double) * (size_t)n);
  double* res2 = (double*)malloc(sizeof(double) * (size_t)n);
  int32_t* i_cell = (int32_t*)calloc((size_t)n, sizeof(int32_t));
  int32_t* first = (int32_t*)malloc(sizeof(int32_t) * (size_t)n);
  for (int64_t i = 0; i < n; ++i)
    for (int k = 0; k < 3; ++k) {
      const double lo = rmin[k] + 0.1 * (rmax[k] - rmin[k]), hi = rmax[k] - 0.1 * (rmax[k] - rmin[k]);
      r[3 * i + k] = lo + uniform01() * (hi - lo);
    }
  CHECK(iu_b200_interpolate_scalar_at_batch(ug, n, r, 1, res, i_cell));     /* cold pass, :46-50 */
  memcpy(first, i_cell, sizeof(int32_t) * (size_t)n);
  CHECK(iu_b200_interpolate_scalar_at_batch(ug, n, r, 1, res2, i_cell));    /* warm pass, :60-64 (dt = 0 on planar meshes) */
  for (int64_t i = 0; i < n; ++i)
    if (first[i] < 1 || i_cell[i] != first[i] || res[i] != res2[i]) { fprintf(stderr, "benchmark flow: pass 2 differs at %lld\n", (long long)i); return 1; }
  int64_t stats[IU_B200_N_STATS];
  CHECK(iu_b200_get_stats(ug, stats));
  if (stats[IU_B200_STAT_KERNEL_LAUNCHES] < 2) { fprintf(stderr, "no kernels launched\n"); return 1; }
  CHECK(iu_b200_grid_destroy(ug));
  free(r); free(res); free(res2); free(i_cell); free(first);
  printf(" benchmark flow PASSED\n");
  return 0;
}

int main(int argc, char** argv) {
  const char* dir = argc > 1 ? argv[1] : "tests/golden";
  if (iu_b200_version() < 100) return 1;
  if (test_interpolation(dir, "triangle", 2, 0)) return 1;
  if (test_interpolation(dir, "quad", 2, 0)) return 1;
  if (test_interpolation(dir, "tetra", 3, 1)) return 1;
  if (benchmark_flow(dir)) return 1;
  printf(" PASSED\n");
  return 0;
}
