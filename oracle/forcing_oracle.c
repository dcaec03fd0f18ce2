/* forcing_oracle.c — plain-C restatement of the reference's forcing-data functions ("port" build of forcing_oracle.h).
 * TEST INFRASTRUCTURE ONLY.  Operation order and comparisons as written in the reference; compiled with -ffp-contract=off. */
#include "forcing_oracle.h"

#include <math.h>

const char *forcing_oracle_kind(void) { return "port"; }

/* src/forcing/rdyforcing_dataset.c:32-66 */
int forcing_oracle_current_data(const double *p, int32_t ndata, double t, int32_t interpolate, int32_t *cur_idx, double *cur) {
  int found = 0;
  double t_up = 0, t_dn = 0, d_up = 0, d_dn = 0;
  for (int32_t i = 0; i < ndata - 1; ++i) {
    t_dn = p[2 * i]; d_dn = p[2 * i + 1]; t_up = p[2 * i + 2]; d_up = p[2 * i + 3];
    if (t >= t_dn && t < t_up) { found = 1; *cur_idx = i; break; }
  }
  if (!found) { *cur_idx = ndata - 1; *cur = p[2 * ndata - 1]; }
  else *cur = interpolate ? (t - t_dn) / (t_up - t_dn) * (d_up - d_dn) + d_dn : d_dn;
  return 0;
}

/* src/forcing/rdyforcing_map.c:78-106 */
int forcing_oracle_map_unstructured(int32_t n, const double *mx, const double *my, int32_t nd, const double *dx_, const double *dy_, int32_t *idx) {
  for (int32_t i = 0; i < n; ++i) {
    double min_dist = 0.0;
    idx[i] = 0;
    for (int32_t k = 0; k < nd; ++k) {
      const double dx = mx[i] - dx_[k], dy = my[i] - dy_[k];
      const double dist = pow(dx * dx + dy * dy, 0.5);
      if (k == 0 || dist < min_dist) { min_dist = dist; idx[i] = k; }
    }
  }
  return 0;
}

/* src/forcing/rdyforcing.c:253-260 (pixel centres, row 0 = northernmost) and src/forcing/rdyforcing_map.c:111-138 */
int forcing_oracle_map_raster(int32_t n, const double *mx, const double *my, int32_t ncols, int32_t nrows, double xlc, double ylc, double cs, int32_t *idx) {
  for (int32_t i = 0; i < n; ++i) {
    double min_dist = ((ncols > nrows ? ncols : nrows) + 1) * cs;
    idx[i] = 0; /* PetscCalloc1: stays 0 when no pixel is closer than the initial bound */
    int32_t k = 0;
    for (int32_t r = 0; r < nrows; ++r) {
      for (int32_t c = 0; c < ncols; ++c, ++k) {
        const double px = xlc + c * cs + cs / 2.0, py = ylc + (nrows - 1 - r) * cs + cs / 2.0;
        const double dx = mx[i] - px, dy = my[i] - py;
        const double dist = pow(dx * dx + dy * dy, 0.5);
        if (dist < min_dist) { min_dist = dist; idx[i] = k; }
      }
    }
  }
  return 0;
}

/* src/forcing/rdyforcing_dataset.c:295-315 */
int forcing_oracle_set_raster(const double *vec, int32_t nvec, int32_t off, int32_t n, const int32_t *idx, double *rain) {
  (void)nvec;
  const double mm_per_hr_2_m_per_sec = 1.0 / (1000.0 * 3600.0);
  for (int32_t i = 0; i < n; ++i) rain[i] = vec[idx[i] + off] * mm_per_hr_2_m_per_sec;
  return 0;
}

/* src/forcing/rdyforcing_dataset.c:350-374 */
int forcing_oracle_set_unstructured(const double *vec, int32_t nvec, int32_t stride, int32_t n, const int32_t *idx, double *values) {
  (void)nvec;
  for (int32_t i = 0; i < n; ++i)
    for (int32_t c = 0; c < stride; ++c) values[i * stride + c] = vec[idx[i] * stride + c + 2];
  return 0;
}

/* src/forcing/rdyforcing_dataset.c:321-344 and :380-406 */
int forcing_oracle_homogeneous(const double *pairs, int32_t ndata, double t, int32_t interpolate, int32_t boundary, int32_t n, double *values, int32_t *cur_idx,
                               int32_t *prev_idx) {
  double v;
  forcing_oracle_current_data(pairs, ndata, t, interpolate, cur_idx, &v);
  if (interpolate || *cur_idx != *prev_idx) {
    *prev_idx = *cur_idx;
    for (int32_t i = 0; i < n; ++i) {
      if (boundary) { values[3 * i] = v; values[3 * i + 1] = 0.0; values[3 * i + 2] = 0.0; }
      else values[i] = v;
    }
  }
  return 0;
}
