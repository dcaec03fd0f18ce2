/* muscl_helpers.h — CPU restatement of the reference's second-order (MUSCL) helpers.  TEST INFRASTRUCTURE ONLY.
 *
 * The reference keeps these four functions in src/operator_fluxes_ceed.c, a libCEED translation unit that cannot be compiled
 * here (it needs the CEED operator API); its caller, ApplyInteriorFlux2R in src/swe/swe_petsc.c:98-213, IS compiled from the
 * reference tree into oracle/_ref.  So for the second-order path the "reference" oracle is the reference's own loop and Roe
 * solver on top of these restated helpers, and the "port" oracle restates the loop as well.  Pin: the second-order MMS
 * convergence thresholds of driver/tests/swe_roe/mms_conv_study_second_order.yaml:57-69 (tests/test_mms.py).
 *
 * Everything works on plain arrays in local numbering:
 *   iedges[nie]        RDyEdges.internal_edge_ids        ecells[2*ne]   RDyEdges.cell_ids
 *   xc, yc [nc]        RDyCells.centroids[].X[0..1]      xm, ym [ne]    edge midpoints 0.5*(v0+v1)
 *   q [3*nc]           the local solution (AoS h, hu, hv)
 * No FMA contraction (compiled with -ffp-contract=off), operation order as written in the reference.
 */
#ifndef MUSCL_HELPERS_H
#define MUSCL_HELPERS_H
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

enum { MUSCL_LIMITER_MINMOD = 0, MUSCL_LIMITER_NONE = 1, MUSCL_LIMITER_VANLEER = 2 }; /* RDyLimiterType, rdyconfigimpl.h:67-71 */

/* PrecomputeLSGradCoeffs, src/operator_fluxes_ceed.c:884-975: coef[4*ie] = {cx_LR, cy_LR, cx_RL, cy_RL} */
static int muscl_ls_coeffs(int32_t nc, int32_t nie, const int32_t *iedges, const int32_t *ecells, const double *xc, const double *yc,
                           double *coef) {
  double *M = calloc(3 * (size_t)(nc > 0 ? nc : 1), sizeof(double)), *inv = calloc(4 * (size_t)(nc > 0 ? nc : 1), sizeof(double));
  if (!M || !inv) { free(M); free(inv); return 55; }
  for (int32_t ie = 0; ie < nie; ++ie) {
    const int32_t e = iedges[ie], cl = ecells[2 * e], cr = ecells[2 * e + 1];
    if (cr < 0) continue;
    const double dx = xc[cr] - xc[cl], dy = yc[cr] - yc[cl];
    const double d = sqrt(dx * dx + dy * dy);
    const double w = (d > 0.0) ? 1.0 / d : 0.0;
    M[cl * 3 + 0] += w * dx * dx; M[cl * 3 + 1] += w * dx * dy; M[cl * 3 + 2] += w * dy * dy;
    M[cr * 3 + 0] += w * dx * dx; M[cr * 3 + 1] += w * dx * dy; M[cr * 3 + 2] += w * dy * dy;
  }
  for (int32_t c = 0; c < nc; ++c) {
    const double m00 = M[c * 3 + 0], m01 = M[c * 3 + 1], m11 = M[c * 3 + 2];
    const double det = m00 * m11 - m01 * m01;
    if (fabs(det) < 1e-15) continue; /* degenerate cell: gradients zeroed */
    const double inv_det = 1.0 / det;
    inv[c * 4 + 0] = m11 * inv_det; inv[c * 4 + 1] = -m01 * inv_det; inv[c * 4 + 2] = -m01 * inv_det; inv[c * 4 + 3] = m00 * inv_det;
  }
  for (int32_t ie = 0; ie < nie; ++ie) {
    const int32_t e = iedges[ie], cl = ecells[2 * e], cr = ecells[2 * e + 1];
    coef[ie * 4 + 0] = coef[ie * 4 + 1] = coef[ie * 4 + 2] = coef[ie * 4 + 3] = 0.0;
    if (cr < 0) continue;
    const double dx = xc[cr] - xc[cl], dy = yc[cr] - yc[cl];
    const double d = sqrt(dx * dx + dy * dy);
    const double w = (d > 0.0) ? 1.0 / d : 0.0;
    const double wdx = w * dx, wdy = w * dy;
    coef[ie * 4 + 0] = inv[cl * 4 + 0] * wdx + inv[cl * 4 + 1] * wdy;
    coef[ie * 4 + 1] = inv[cl * 4 + 2] * wdx + inv[cl * 4 + 3] * wdy;
    coef[ie * 4 + 2] = inv[cr * 4 + 0] * wdx + inv[cr * 4 + 1] * wdy;
    coef[ie * 4 + 3] = inv[cr * 4 + 2] * wdx + inv[cr * 4 + 3] * wdy;
  }
  free(M); free(inv);
  return 0;
}

/* ComputeLeastSquaresGradients, src/operator_fluxes_ceed.c:993-1038: g*[2*c + {0,1}] = d/dx, d/dy */
static void muscl_gradients(int32_t nc, int32_t nie, const int32_t *iedges, const int32_t *ecells, const double *coef, const double *q,
                            double *gh, double *ghu, double *ghv) {
  for (int32_t i = 0; i < 2 * nc; ++i) gh[i] = ghu[i] = ghv[i] = 0.0;
  for (int32_t ie = 0; ie < nie; ++ie) {
    const int32_t e = iedges[ie], cl = ecells[2 * e], cr = ecells[2 * e + 1];
    if (cr < 0) continue;
    const double cx_lr = coef[ie * 4 + 0], cy_lr = coef[ie * 4 + 1], cx_rl = coef[ie * 4 + 2], cy_rl = coef[ie * 4 + 3];
    const double dq_h = q[cr * 3 + 0] - q[cl * 3 + 0], dq_hu = q[cr * 3 + 1] - q[cl * 3 + 1], dq_hv = q[cr * 3 + 2] - q[cl * 3 + 2];
    gh[cl * 2 + 0] += cx_lr * dq_h;   gh[cl * 2 + 1] += cy_lr * dq_h;
    ghu[cl * 2 + 0] += cx_lr * dq_hu; ghu[cl * 2 + 1] += cy_lr * dq_hu;
    ghv[cl * 2 + 0] += cx_lr * dq_hv; ghv[cl * 2 + 1] += cy_lr * dq_hv;
    gh[cr * 2 + 0] += cx_rl * dq_h;   gh[cr * 2 + 1] += cy_rl * dq_h;
    ghu[cr * 2 + 0] += cx_rl * dq_hu; ghu[cr * 2 + 1] += cy_rl * dq_hu;
    ghv[cr * 2 + 0] += cx_rl * dq_hv; ghv[cr * 2 + 1] += cy_rl * dq_hv;
  }
}

/* Minmod / VanLeer / LimitSlope, src/operator_fluxes_ceed.c:1100-1132 */
static double muscl_limit(int limiter, double extrap, double half_dq) {
  if (limiter == MUSCL_LIMITER_NONE) return extrap;
  if (extrap * half_dq <= 0.0) return 0.0;
  if (limiter == MUSCL_LIMITER_VANLEER) return 2.0 * extrap * half_dq / (extrap + half_dq);
  return fabs(extrap) < fabs(half_dq) ? extrap : half_dq;
}

/* ReconstructFaceValues, src/operator_fluxes_ceed.c:1150-1195: qface[6*k] = {h_L, hu_L, hv_L, h_R, hu_R, hv_R} for the k-th
 * OWNED internal edge (edge_owned == NULL: every edge is owned) */
static void muscl_faces(int32_t nie, const int32_t *iedges, const int32_t *ecells, const int32_t *edge_owned, const double *xc,
                        const double *yc, const double *xm, const double *ym, const double *q, const double *gh, const double *ghu,
                        const double *ghv, int limiter, double *qface) {
  int32_t k = 0;
  for (int32_t ie = 0; ie < nie; ++ie) {
    const int32_t e = iedges[ie];
    if (edge_owned && !edge_owned[e]) continue;
    const int32_t cl = ecells[2 * e], cr = ecells[2 * e + 1];
    if (cr < 0) { for (int c = 0; c < 6; ++c) qface[k * 6 + c] = 0.0; ++k; continue; }
    const double dx_l = xm[e] - xc[cl], dy_l = ym[e] - yc[cl], dx_r = xm[e] - xc[cr], dy_r = ym[e] - yc[cr];
    const double *g[3] = {gh, ghu, ghv};
    for (int c = 0; c < 3; ++c) {
      const double extrap_l = g[c][cl * 2 + 0] * dx_l + g[c][cl * 2 + 1] * dy_l;
      const double extrap_r = g[c][cr * 2 + 0] * dx_r + g[c][cr * 2 + 1] * dy_r;
      const double dq = q[cr * 3 + c] - q[cl * 3 + c];
      qface[k * 6 + c]     = q[cl * 3 + c] + muscl_limit(limiter, extrap_l, 0.5 * dq);
      qface[k * 6 + 3 + c] = q[cr * 3 + c] + muscl_limit(limiter, extrap_r, -0.5 * dq);
    }
    qface[k * 6 + 0] = fmax(0.0, qface[k * 6 + 0]);
    qface[k * 6 + 3] = fmax(0.0, qface[k * 6 + 3]);
    ++k;
  }
}
#endif
