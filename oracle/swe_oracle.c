/* swe_oracle.c — CPU restatement ("port") of RDycore's PETSc-backend SWE operator.
 * TEST INFRASTRUCTURE ONLY: see swe_oracle.h for who may load it.
 *
 * Written from scratch following the reference's arithmetic statement by statement so that,
 * compiled with -ffp-contract=off, it reproduces oracle/_ref (the reference's own code) bit for
 * bit; tests/test_oracle.py asserts that.  Parity status: PINNED against oracle/_ref for every
 * kernel below; the TS stage update (PETSc TSEULER/TSRK4) and DMPlex geometry are third-party code
 * absent from /root/reference and stay "parity unpinned" (DESIGN.md).
 *
 * Reference locations (relative to the RDycore tree):
 *   velocities                     src/swe/swe_petsc.c:57-73
 *   Roe eigenspectrum / flux       src/swe/swe_roe_flux_petsc.h:15-81, 91-132
 *   interior edges (1st order)     src/swe/swe_petsc.c:215-316
 *   boundary edges + 3 BC kinds    src/swe/swe_petsc.c:434-630
 *   sources: semi-implicit/XQ2018  src/swe/swe_petsc.c:704-804, 816-932
 *   HR interior edges              src/swe/swe_petsc.c:1000-1161
 *   operator composition           src/operator.c:656-672, src/operator_fluxes_petsc.c:28-51
 *   RHS callback                   src/rdysetup.c:1120-1139
 */
#include "swe_oracle.h"
#include "muscl_helpers.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static const double G = 9.806; /* src/swe/swe_types_petsc.h:7 */

typedef struct {
  int32_t  ne, type;
  int32_t *edge_ids;
  double  *values, *fluxes, *accum;  /* [3*ne] AoS */
  double  *hl, *ul, *vl, *hr, *ur, *vr, *amax; /* scratch persists between calls like the reference's (swe_petsc.c:7-43) */
} OBoundary;

struct SweOracle_ {
  RdybConfig cfg;
  int32_t    nc, nown, ne, nie, nb;
  int32_t   *is_owned, *l2o, *ecells, *iedges;
  int64_t   *cgid, *egid;
  double    *area, *dzdx, *dzdy, *zc, *len, *cn, *sn;
  OBoundary *bnd;
  double    *u_global, *u_local, *rhs, *ext, *mat, *fdiv, *prim, *src_inst;
  double    *k[4], *ystage;
  RdybCourantDiagnostics courant;
  /* second order (MUSCL): limiter (-1 = first order), geometry, LS weights, work arrays */
  int     limiter;
  double *xc, *yc, *xm, *ym, *ls_coef, *gh, *ghu, *ghv, *qface;
};

const char *swe_oracle_kind(void) { return "port"; }

static void *dup_arr(const void *src, size_t n, size_t sz) {
  void *p = calloc(n > 0 ? n : 1, sz);
  if (src && n) memcpy(p, src, n * sz);
  return p;
}

int swe_oracle_create(const RdybConfig *cfg, const RdybMeshView *mv, int32_t nb, const RdybBoundaryView *bv, SweOracle *out) {
  SweOracle o = calloc(1, sizeof(*o));
  *out = o;
  o->cfg = *cfg;
  o->nc = mv->num_cells; o->nown = mv->num_owned_cells; o->ne = mv->num_edges; o->nie = mv->num_internal_edges; o->nb = nb;
  o->is_owned = dup_arr(mv->cell_is_owned, (size_t)o->nc, sizeof(int32_t));
  o->l2o      = dup_arr(mv->cell_local_to_owned, (size_t)o->nc, sizeof(int32_t));
  o->ecells   = dup_arr(mv->edge_cell_ids, 2 * (size_t)o->ne, sizeof(int32_t));
  o->iedges   = dup_arr(mv->internal_edge_ids, (size_t)o->nie, sizeof(int32_t));
  o->cgid     = dup_arr(mv->cell_global_ids, (size_t)o->nc, sizeof(int64_t));
  o->egid     = dup_arr(mv->edge_global_ids, (size_t)o->ne, sizeof(int64_t));
  if (!mv->cell_global_ids) for (int32_t c = 0; c < o->nc; ++c) o->cgid[c] = c;
  if (!mv->edge_global_ids) for (int32_t e = 0; e < o->ne; ++e) o->egid[e] = e;
  o->area = dup_arr(mv->cell_areas, (size_t)o->nc, sizeof(double));
  o->dzdx = dup_arr(mv->cell_dz_dx, (size_t)o->nc, sizeof(double));
  o->dzdy = dup_arr(mv->cell_dz_dy, (size_t)o->nc, sizeof(double));
  o->zc   = dup_arr(mv->cell_zc, (size_t)o->nc, sizeof(double));
  o->len  = dup_arr(mv->edge_lengths, (size_t)o->ne, sizeof(double));
  o->cn   = dup_arr(mv->edge_cn, (size_t)o->ne, sizeof(double));
  o->sn   = dup_arr(mv->edge_sn, (size_t)o->ne, sizeof(double));
  if (cfg->well_balancing == RDYB_WELL_BALANCING_HR && !mv->cell_zc) { fprintf(stderr, "swe_oracle: HR needs cell_zc\n"); return RDYB_ERR_USER; }
  if (cfg->source_method != RDYB_SOURCE_SEMI_IMPLICIT && cfg->source_method != RDYB_SOURCE_IMPLICIT_XQ2018) {
    /* swe_petsc.c:973-975 */
    fprintf(stderr, "Only semi_implicit and implicit_xq2018 are supported in the PETSc version\n");
    return RDYB_ERR_USER;
  }
  /* numerics.second_order: only the first-order (non-HR) interior operator has the MUSCL variant (swe_petsc.c:357-376, 404);
   * the combination with HR is refused at operator creation (src/operator.c:388-389) */
  o->limiter = -1;
  if (cfg->reconstruction != RDYB_RECON_FIRST_ORDER && cfg->well_balancing == RDYB_WELL_BALANCING_HR) {
    fprintf(stderr, "-second_order cannot be used with well_balancing = HR simultaneously (not yet implemented)\n");
    return RDYB_ERR_USER;
  }
  if (cfg->reconstruction != RDYB_RECON_FIRST_ORDER) {
    if (!mv->cell_xc || !mv->cell_yc || !mv->edge_xm || !mv->edge_ym) { fprintf(stderr, "swe_oracle: second order needs centroids and edge midpoints\n"); return RDYB_ERR_USER; }
    o->limiter = cfg->reconstruction == RDYB_RECON_MUSCL_NONE ? MUSCL_LIMITER_NONE : cfg->reconstruction == RDYB_RECON_MUSCL_VANLEER ? MUSCL_LIMITER_VANLEER : MUSCL_LIMITER_MINMOD;
    o->xc = dup_arr(mv->cell_xc, (size_t)o->nc, 8); o->yc = dup_arr(mv->cell_yc, (size_t)o->nc, 8);
    o->xm = dup_arr(mv->edge_xm, (size_t)o->ne, 8); o->ym = dup_arr(mv->edge_ym, (size_t)o->ne, 8);
    o->ls_coef = dup_arr(NULL, 4 * (size_t)o->nie, 8); o->qface = dup_arr(NULL, 6 * (size_t)o->nie, 8);
    o->gh = dup_arr(NULL, 2 * (size_t)o->nc, 8); o->ghu = dup_arr(NULL, 2 * (size_t)o->nc, 8); o->ghv = dup_arr(NULL, 2 * (size_t)o->nc, 8);
    if (muscl_ls_coeffs(o->nc, o->nie, o->iedges, o->ecells, o->xc, o->yc, o->ls_coef)) return RDYB_ERR_MEM;
  }
  o->bnd = calloc((size_t)(nb > 0 ? nb : 1), sizeof(OBoundary));
  for (int32_t b = 0; b < nb; ++b) {
    OBoundary *B = &o->bnd[b];
    size_t n = (size_t)bv[b].num_edges;
    B->ne = bv[b].num_edges; B->type = bv[b].bc_type;
    B->edge_ids = dup_arr(bv[b].edge_ids, n, sizeof(int32_t));
    B->values = dup_arr(NULL, 3 * n, sizeof(double)); B->fluxes = dup_arr(NULL, 3 * n, sizeof(double)); B->accum = dup_arr(NULL, 3 * n, sizeof(double));
    B->hl = dup_arr(NULL, n, 8); B->ul = dup_arr(NULL, n, 8); B->vl = dup_arr(NULL, n, 8); B->hr = dup_arr(NULL, n, 8);
    B->ur = dup_arr(NULL, n, 8); B->vr = dup_arr(NULL, n, 8); B->amax = dup_arr(NULL, n, 8);
  }
  size_t N = (size_t)o->nown;
  o->u_global = dup_arr(NULL, 3 * N, 8); o->rhs = dup_arr(NULL, 3 * N, 8); o->ext = dup_arr(NULL, 3 * N, 8); o->mat = dup_arr(NULL, N, 8);
  o->fdiv = dup_arr(NULL, 3 * N, 8); o->prim = dup_arr(NULL, 3 * N, 8); o->src_inst = dup_arr(NULL, 3 * N, 8);
  o->u_local = dup_arr(NULL, 3 * (size_t)o->nc, 8); o->ystage = dup_arr(NULL, 3 * N, 8);
  for (int i = 0; i < 4; ++i) o->k[i] = dup_arr(NULL, 3 * N, 8);
  return 0;
}

void swe_oracle_destroy(SweOracle o) {
  if (!o) return;
  for (int32_t b = 0; b < o->nb; ++b) {
    OBoundary *B = &o->bnd[b];
    free(B->edge_ids); free(B->values); free(B->fluxes); free(B->accum);
    free(B->hl); free(B->ul); free(B->vl); free(B->hr); free(B->ur); free(B->vr); free(B->amax);
  }
  free(o->bnd); free(o->is_owned); free(o->l2o); free(o->ecells); free(o->iedges); free(o->cgid); free(o->egid);
  free(o->area); free(o->dzdx); free(o->dzdy); free(o->zc); free(o->len); free(o->cn); free(o->sn);
  free(o->u_global); free(o->u_local); free(o->rhs); free(o->ext); free(o->mat); free(o->fdiv); free(o->prim); free(o->src_inst);
  free(o->ystage);
  for (int i = 0; i < 4; ++i) free(o->k[i]);
  free(o->xc); free(o->yc); free(o->xm); free(o->ym); free(o->ls_coef); free(o->gh); free(o->ghu); free(o->ghv); free(o->qface);
  free(o);
}

double *swe_oracle_array(SweOracle o, const char *name, int32_t b, int64_t *length) {
  double *p = NULL;
  int64_t n = 0, N = o->nown;
  if (!strcmp(name, "u_global")) { p = o->u_global; n = 3 * N; }
  else if (!strcmp(name, "u_local")) { p = o->u_local; n = 3 * (int64_t)o->nc; }
  else if (!strcmp(name, "rhs")) { p = o->rhs; n = 3 * N; }
  else if (!strcmp(name, "external_sources")) { p = o->ext; n = 3 * N; }
  else if (!strcmp(name, "material_properties")) { p = o->mat; n = N; }
  else if (!strcmp(name, "flux_divergence")) { p = o->fdiv; n = 3 * N; }
  else if (!strcmp(name, "primitive_variables")) { p = o->prim; n = 3 * N; }
  else if (!strcmp(name, "src_inst")) { p = o->src_inst; n = 3 * N; }
  else if (b >= 0 && b < o->nb) {
    n = 3 * (int64_t)o->bnd[b].ne;
    if (!strcmp(name, "boundary_values")) p = o->bnd[b].values;
    else if (!strcmp(name, "boundary_fluxes")) p = o->bnd[b].fluxes;
    else if (!strcmp(name, "boundary_fluxes_accum")) p = o->bnd[b].accum;
  }
  if (p && length) *length = n;
  return p;
}

/* ---------------- arithmetic ---------------- */
/* swe_petsc.c:57-73 — wet test is h < tiny_h => dry */
static void velocities(double tiny_h, double ha, double h, double hu, double hv, double *u, double *v) {
  if (h < tiny_h) { *u = 0.0; *v = 0.0; }
  else {
    double denom = h * h + ha * ha;
    *u = hu * h / denom;
    *v = hv * h / denom;
  }
}

/* swe_roe_flux_petsc.h:15-81 and :108-128, one edge.  pow(x, 0.5) is kept (not sqrt). */
static void roe_flux(double hl, double ul, double vl, double hr, double ur, double vr, double sn, double cn, double f[3], double *amax) {
  double duml = pow(hl, 0.5), dumr = pow(hr, 0.5);
  double cl = pow(G * hl, 0.5), cr = pow(G * hr, 0.5);
  double hhat = duml * dumr;
  double uhat = (duml * ul + dumr * ur) / (duml + dumr);
  double vhat = (duml * vl + dumr * vr) / (duml + dumr);
  double chat = pow(0.5 * G * (hl + hr), 0.5);
  double uperp = uhat * cn + vhat * sn;
  double dh = hr - hl, du = ur - ul, dv = vr - vl;
  double dupar = -du * sn + dv * cn;
  double duperp = du * cn + dv * sn;
  double R[3][3] = {{1.0, 0.0, 1.0}, {uhat - chat * cn, -sn, uhat + chat * cn}, {vhat - chat * sn, cn, vhat + chat * sn}};
  double uperpl = ul * cn + vl * sn, uperpr = ur * cn + vr * sn;
  double a1 = fabs(uperp - chat), a2 = fabs(uperp), a3 = fabs(uperp + chat);
  /* critical flow fix */
  double al1 = uperpl - cl, ar1 = uperpr - cr;
  double da1 = fmax(0.0, 2.0 * (ar1 - al1));
  if (a1 < da1) a1 = 0.5 * (a1 * a1 / da1 + da1);
  double al3 = uperpl + cl, ar3 = uperpr + cr;
  double da3 = fmax(0.0, 2.0 * (ar3 - al3));
  if (a3 < da3) a3 = 0.5 * (a3 * a3 / da3 + da3);
  double A[3] = {a1, a2, a3};
  double dW[3] = {0.5 * (dh - hhat * duperp / chat), hhat * dupar, 0.5 * (dh + hhat * duperp / chat)};
  *amax = chat + fabs(uperp);
  double FL[3] = {uperpl * hl, ul * uperpl * hl + 0.5 * G * hl * hl * cn, vl * uperpl * hl + 0.5 * G * hl * hl * sn};
  double FR[3] = {uperpr * hr, ur * uperpr * hr + 0.5 * G * hr * hr * cn, vr * uperpr * hr + 0.5 * G * hr * hr * sn};
  for (int i = 0; i < 3; ++i) f[i] = 0.5 * (FL[i] + FR[i] - R[i][0] * A[0] * dW[0] - R[i][1] * A[1] * dW[1] - R[i][2] * A[2] * dW[2]);
}

static void courant_update(SweOracle o, double cnum, int64_t egid, int64_t cgid) {
  if (cnum > o->courant.max_courant_num) {
    o->courant.max_courant_num = cnum;
    o->courant.global_edge_id = egid;
    o->courant.global_cell_id = cgid;
  }
}

/* swe_petsc.c:215-316 */
static void interior_flux(SweOracle o, double dt, const double *u, double *F) {
  const double tiny = o->cfg.tiny_h, ha = o->cfg.h_anuga_regular;
  for (int32_t i = 0; i < o->nie; ++i) {
    int32_t e = o->iedges[i], l = o->ecells[2 * e], r = o->ecells[2 * e + 1];
    if (r == -1) continue;
    double hl = u[3 * l], hr = u[3 * r], ul, vl, ur, vr;
    if (hr < tiny && hl < tiny) continue; /* the reference computes then discards this edge (:266 vs :285) */
    velocities(tiny, ha, hl, u[3 * l + 1], u[3 * l + 2], &ul, &vl);
    velocities(tiny, ha, hr, u[3 * r + 1], u[3 * r + 2], &ur, &vr);
    double f[3], amax;
    roe_flux(hl, ul, vl, hr, ur, vr, o->sn[e], o->cn[e], f, &amax);
    double len = o->len[e], al = o->area[l], ar = o->area[r];
    double cnum = amax * len / fmin(al, ar) * dt;
    courant_update(o, cnum, o->egid[e], al < ar ? o->cgid[l] : o->cgid[r]);
    for (int d = 0; d < 3; ++d) {
      if (o->is_owned[l]) F[3 * o->l2o[l] + d] += f[d] * (-len / al);
      if (o->is_owned[r]) F[3 * o->l2o[r] + d] += f[d] * (len / ar);
    }
  }
}

/* ApplyInteriorFlux2R, swe_petsc.c:98-213 (serial: every edge is owned, DMLocalToGlobal(ADD) adds the local vector's owned rows
 * to the zeroed F; CommunicateCellGradients is the identity) */
static void interior_flux_2r(SweOracle o, double dt, const double *u, double *F) {
  const double tiny = o->cfg.tiny_h, ha = o->cfg.h_anuga_regular;
  muscl_gradients(o->nc, o->nie, o->iedges, o->ecells, o->ls_coef, u, o->gh, o->ghu, o->ghv);
  muscl_faces(o->nie, o->iedges, o->ecells, NULL, o->xc, o->yc, o->xm, o->ym, u, o->gh, o->ghu, o->ghv, o->limiter, o->qface);
  double *loc = calloc(3 * (size_t)(o->nc > 0 ? o->nc : 1), sizeof(double)); /* rhs_local */
  for (int32_t i = 0; i < o->nie; ++i) {
    int32_t e = o->iedges[i], l = o->ecells[2 * e], r = o->ecells[2 * e + 1];
    if (r < 0) continue;
    const double *qf = &o->qface[6 * (size_t)i];
    double hl = fmax(0.0, qf[0]), hr = fmax(0.0, qf[3]), ul, vl, ur, vr;
    if (hr < tiny && hl < tiny) continue; /* computed then discarded by the reference (:155 vs :179) */
    velocities(tiny, ha, hl, qf[1], qf[2], &ul, &vl);
    velocities(tiny, ha, hr, qf[4], qf[5], &ur, &vr);
    double f[3], amax;
    roe_flux(hl, ul, vl, hr, ur, vr, o->sn[e], o->cn[e], f, &amax);
    double len = o->len[e], al = o->area[l], ar = o->area[r];
    double cnum = amax * len / fmin(al, ar) * dt;
    courant_update(o, cnum, o->egid[e], al < ar ? o->cgid[l] : o->cgid[r]);
    for (int d = 0; d < 3; ++d) {
      loc[3 * l + d] += f[d] * (-len / al);
      loc[3 * r + d] += f[d] * (len / ar);
    }
  }
  for (int32_t c = 0; c < o->nc; ++c)
    if (o->is_owned[c])
      for (int d = 0; d < 3; ++d) F[3 * o->l2o[c] + d] += loc[3 * c + d];
  free(loc);
}

/* swe_petsc.c:1000-1161 */
static void interior_flux_hr(SweOracle o, double dt, const double *u, double *F) {
  const double tiny = o->cfg.tiny_h, ha = o->cfg.h_anuga_regular;
  for (int32_t i = 0; i < o->nie; ++i) {
    int32_t e = o->iedges[i], l = o->ecells[2 * e], r = o->ecells[2 * e + 1];
    if (r == -1) continue;
    double h_L = u[3 * l], hu_L = u[3 * l + 1], hv_L = u[3 * l + 2], h_R = u[3 * r], hu_R = u[3 * r + 1], hv_R = u[3 * r + 2];
    if (h_R < tiny && h_L < tiny) continue;
    double zL = o->zc[l], zR = o->zc[r];
    double etaL = h_L + zL, etaR = h_R + zR, zmax = fmax(zL, zR);
    double hLs = fmax(0.0, etaL - zmax), hRs = fmax(0.0, etaR - zmax);
    double dL = h_L * h_L + ha * ha, dR = h_R * h_R + ha * ha;
    /* strict > here, unlike velocities() */
    double uL = (h_L > tiny) ? hu_L * h_L / dL : 0.0, vL = (h_L > tiny) ? hv_L * h_L / dL : 0.0;
    double uR = (h_R > tiny) ? hu_R * h_R / dR : 0.0, vR = (h_R > tiny) ? hv_R * h_R / dR : 0.0;
    double len = o->len[e], al = o->area[l], ar = o->area[r], cn = o->cn[e], sn = o->sn[e];
    double sl = -len / al, sr = len / ar;
    if (hLs > tiny || hRs > tiny) {
      double f[3], amax;
      roe_flux(hLs, uL, vL, hRs, uR, vR, sn, cn, f, &amax);
      double cnum = amax * len / fmin(al, ar) * dt;
      courant_update(o, cnum, o->egid[e], al < ar ? o->cgid[l] : o->cgid[r]);
      for (int d = 0; d < 3; ++d) {
        if (o->is_owned[l]) F[3 * o->l2o[l] + d] += f[d] * sl;
        if (o->is_owned[r]) F[3 * o->l2o[r] + d] += f[d] * sr;
      }
    }
    double corrL = 0.5 * G * (h_L * h_L - hLs * hLs), corrR = 0.5 * G * (h_R * h_R - hRs * hRs);
    if (o->is_owned[l]) { F[3 * o->l2o[l] + 1] += corrL * cn * sl; F[3 * o->l2o[l] + 2] += corrL * sn * sl; }
    if (o->is_owned[r]) { F[3 * o->l2o[r] + 1] += corrR * cn * sr; F[3 * o->l2o[r] + 2] += corrR * sn * sr; }
  }
}

/* swe_petsc.c:434-630 */
static void boundary_flux(SweOracle o, OBoundary *B, double dt, const double *u, double *F) {
  const double tiny = o->cfg.tiny_h, ha = o->cfg.h_anuga_regular;
  for (int32_t k = 0; k < B->ne; ++k) {
    int32_t e = B->edge_ids[k], l = o->ecells[2 * e];
    double sn = o->sn[e], cn = o->cn[e];
    B->hl[k] = u[3 * l];
    velocities(tiny, ha, B->hl[k], u[3 * l + 1], u[3 * l + 2], &B->ul[k], &B->vl[k]);
    switch (B->type) {
      case RDYB_CONDITION_DIRICHLET:
        B->hr[k] = B->values[3 * k];
        velocities(tiny, ha, B->hr[k], B->values[3 * k + 1], B->values[3 * k + 2], &B->ur[k], &B->vr[k]);
        break;
      case RDYB_CONDITION_REFLECTING:
        if (o->is_owned[l]) {
          double dum1 = sn * sn - cn * cn, dum2 = 2.0 * sn * cn;
          B->hr[k] = B->hl[k];
          B->ur[k] = B->ul[k] * dum1 - B->vl[k] * dum2;
          B->vr[k] = -B->ul[k] * dum2 - B->vl[k] * dum1;
        }
        break;
      case RDYB_CONDITION_CRITICAL_OUTFLOW:
        if (o->is_owned[l]) {
          double uperp = B->ul[k] * cn + B->vl[k] * sn;
          if (uperp < 0.0) {
            B->hl[k] = B->ul[k] = B->vl[k] = 0.0;
            B->hr[k] = B->ur[k] = B->vr[k] = 0.0;
          } else {
            double q = B->hl[k] * fabs(uperp);
            B->hr[k] = pow(q * q / G, 1.0 / 3.0);
            double vel = pow(G * B->hr[k], 0.5);
            B->ur[k] = vel * cn;
            B->vr[k] = vel * sn;
          }
        }
        break;
      default:
        break;
    }
    /* Roe runs before the wet/dry guard (:574 vs :593): dry-dry edges leave NaN in boundary_fluxes, as in the reference */
    roe_flux(B->hl[k], B->ul[k], B->vl[k], B->hr[k], B->ur[k], B->vr[k], sn, cn, &B->fluxes[3 * k], &B->amax[k]);
    if (o->is_owned[l] && !(B->hl[k] < tiny && B->hr[k] < tiny)) {
      double len = o->len[e], a = o->area[l];
      double cnum = B->amax[k] * len / a * dt;
      courant_update(o, cnum, o->egid[e], o->cgid[l]);
      for (int d = 0; d < 3; ++d) F[3 * o->l2o[l] + d] += B->fluxes[3 * k + d] * (-len / a);
    }
  }
  for (int32_t i = 0; i < 3 * B->ne; ++i) B->accum[i] += dt * B->fluxes[i]; /* VecAXPY, :623 */
}

/* swe_petsc.c:704-804 and 816-932 */
static void sources(SweOracle o, double dt, const double *u, double *F) {
  const double tiny = o->cfg.tiny_h, ha = o->cfg.h_anuga_regular, thr = o->cfg.xq2018_threshold;
  const int bed = o->cfg.well_balancing != RDYB_WELL_BALANCING_HR;
  const int xq = o->cfg.source_method == RDYB_SOURCE_IMPLICIT_XQ2018;
  for (int32_t c = 0; c < o->nc; ++c) {
    if (!o->is_owned[c]) continue;
    int32_t oc = o->l2o[c];
    double h = u[3 * c], hu = u[3 * c + 1], hv = u[3 * c + 2];
    double bedx = 0.0, bedy = 0.0, tbx = 0.0, tby = 0.0;
    if (bed) { bedx = o->dzdx[c] * G * h; bedy = o->dzdy[c] * G * h; }
    if (h >= tiny) {
      double n = o->mat[oc], Fx = o->fdiv[3 * oc + 1], Fy = o->fdiv[3 * oc + 2];
      if (!xq) {
        double uu = hu / h, vv = hv / h;
        double Cd = G * (n * n) * pow(h, -1.0 / 3.0);
        double vel = sqrt(uu * uu + vv * vv);
        double tb = Cd * vel / h;
        double factor = tb / (1.0 + dt * tb);
        tbx = (hu + dt * Fx - dt * bedx) * factor;
        tby = (hv + dt * Fy - dt * bedy) * factor;
      } else {
        double Ax = Fx - bedx, Ay = Fy - bedy;
        double mx = hu + Ax * dt, my = hv + Ay * dt;
        double lambda = G * (n * n) * pow(h, -4.0 / 3.0) * pow((mx / h) * (mx / h) + (my / h) * (my / h), 0.5);
        double qx, qy;
        if (dt * lambda < thr) { qx = mx; qy = my; }
        else {
          qx = (mx - mx * pow(1.0 + 4.0 * dt * lambda, 0.5)) / (-2.0 * dt * lambda);
          qy = (my - my * pow(1.0 + 4.0 * dt * lambda, 0.5)) / (-2.0 * dt * lambda);
        }
        double qmag = pow(qx * qx + qy * qy, 0.5);
        tbx = G * (n * n) * pow(h, -7.0 / 3.0) * qx * qmag;
        tby = G * (n * n) * pow(h, -7.0 / 3.0) * qy * qmag;
      }
    }
    F[3 * oc] += o->ext[3 * oc];
    F[3 * oc + 1] += -bedx - tbx + o->ext[3 * oc + 1];
    F[3 * oc + 2] += -bedy - tby + o->ext[3 * oc + 2];
    double denom = h * h + ha * ha;
    o->prim[3 * oc] = h;
    o->prim[3 * oc + 1] = (h >= tiny) ? (hu * h / denom) : 0.0;
    o->prim[3 * oc + 2] = (h >= tiny) ? (hv * h / denom) : 0.0;
  }
}

/* src/rdysetup.c:1127-1139 + src/operator.c:656-672 */
static int rhs_function(SweOracle o, double dt, const double *U, double *F) {
  size_t N = (size_t)o->nown;
  memset(F, 0, 3 * N * sizeof(double));
  for (int32_t c = 0; c < o->nc; ++c)
    if (o->is_owned[c]) memcpy(&o->u_local[3 * c], &U[3 * (size_t)o->l2o[c]], 3 * sizeof(double));
  o->courant.max_courant_num = 0.0; o->courant.global_edge_id = -1; o->courant.global_cell_id = -1;
  if (o->cfg.well_balancing == RDYB_WELL_BALANCING_HR) interior_flux_hr(o, dt, o->u_local, F);
  else if (o->limiter >= 0) interior_flux_2r(o, dt, o->u_local, F);
  else interior_flux(o, dt, o->u_local, F);
  for (int32_t b = 0; b < o->nb; ++b) boundary_flux(o, &o->bnd[b], dt, o->u_local, F);
  memcpy(o->fdiv, F, 3 * N * sizeof(double));
  sources(o, dt, o->u_local, F);
  memcpy(o->src_inst, o->ext, 3 * N * sizeof(double));
  return 0;
}

int swe_oracle_rhs(SweOracle o, double dt) { return rhs_function(o, dt, o->u_global, o->rhs); }

int swe_oracle_euler_update(SweOracle o, double dt) {
  for (int64_t i = 0; i < 3 * (int64_t)o->nown; ++i) o->u_global[i] = o->u_global[i] + dt * o->rhs[i];
  return 0;
}

int swe_oracle_step_euler(SweOracle o, double dt, int32_t nsteps) {
  for (int32_t s = 0; s < nsteps; ++s) { swe_oracle_rhs(o, dt); swe_oracle_euler_update(o, dt); }
  return 0;
}

int swe_oracle_step_rk4(SweOracle o, double dt, int32_t nsteps) {
  static const double A[4][3] = {{0, 0, 0}, {0.5, 0, 0}, {0, 0.5, 0}, {0, 0, 1.0}};
  static const double bw[4] = {1.0 / 6.0, 1.0 / 3.0, 1.0 / 3.0, 1.0 / 6.0};
  int64_t n = 3 * (int64_t)o->nown;
  for (int32_t s = 0; s < nsteps; ++s) {
    for (int i = 0; i < 4; ++i) {
      memcpy(o->ystage, o->u_global, (size_t)n * sizeof(double));
      for (int j = 0; j < i; ++j) {
        double w = dt * A[i][j];
        if (w != 0.0) for (int64_t q = 0; q < n; ++q) o->ystage[q] += w * o->k[j][q];
      }
      rhs_function(o, dt, o->ystage, o->k[i]);
    }
    for (int j = 0; j < 4; ++j) {
      double w = dt * bw[j];
      for (int64_t q = 0; q < n; ++q) o->u_global[q] += w * o->k[j][q];
    }
  }
  return 0;
}

int swe_oracle_courant(SweOracle o, RdybCourantDiagnostics *d) { *d = o->courant; return 0; }
