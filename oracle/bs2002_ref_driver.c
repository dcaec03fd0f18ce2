/* bs2002_ref_driver.c — CPU oracle of the libCEED-semantics SWE operator with Bradford-Sanders (2002) well balancing.
 * TEST INFRASTRUCTURE ONLY (see swe_oracle.h).
 *
 * `physics.flow.well_balancing: bradford_sanders_2002` exists only in the reference's libCEED backend (src/operator.c:170-192; the
 * PETSc backend runs such decks on the standard operators, :202-209).  Its arithmetic lives in four Q-function headers, which are
 * #included here IN PLACE from /root/reference behind oracle/ceed_qf_shim (a 10-line <ceed/types.h>):
 *   src/swe/swe_well_balancing.h  SWEEtaVertex             water-surface elevation of a cell as seen by one of its vertices
 *   src/swe/swe_fluxes_ceed.h     SWEFlux_Roe, SWEBoundaryFlux_{Dirichlet,Reflecting,Outflow}_Roe  (ComputeDhv feeding
 *   src/swe/swe_roe_flux_ceed.h   SWERiemannFlux_Roe        flux[1,2] += g dhv^2 / 24 (cn, sn))
 *   src/swe/swe_sources_ceed.h    SWESourcesWith{SemiImplicitBedFriction,ImplicitBedFrictionXQ2018}
 * so every floating-point operation of the flux, the well-balancing term and the sources is the reference's own code.  What libCEED
 * itself does - gathering Q-function inputs through element restrictions and summing outputs through their transposes - is
 * restated below from the operator constructors (src/operator_fluxes_ceed.c:150-260, 470-560, 656-859; src/operator_sources_ceed.c)
 * with the CPU /ref backend's order: one output field after the other, elements in ascending order.
 * Serial meshes only (the well-balancing row of SURVEY 8f-4 is a single-GPU row).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <swe_well_balancing.h>
#include <swe_fluxes_ceed.h>
#include <swe_sources_ceed.h>

typedef struct {
  int32_t num_cells, num_vertices, num_edges, num_internal_edges, num_boundary_edges;
  const int32_t *cell_vertex_offsets, *cell_vertex_ids; /* [num_cells+1], triangles */
  const double  *vertex_z;                              /* [num_vertices] */
  const int32_t *edge_cell_ids, *edge_vertex_ids;       /* [2*num_edges] */
  const double  *edge_lengths, *edge_cn, *edge_sn;
  const int32_t *internal_edge_ids;
  const double  *cell_areas, *cell_dz_dx, *cell_dz_dy;
  /* boundary edges, flattened over the boundaries in boundary order */
  const int32_t *boundary_edge_ids, *boundary_edge_type; /* [num_boundary_edges]: RDyConditionType 0 Dirichlet, 2 reflecting, 3 critical outflow */
  const double  *boundary_values;                        /* [num_boundary_edges][3] Dirichlet states (ignored otherwise) */
} Bs2002Mesh;

typedef struct {
  double  tiny_h, h_anuga_regular, xq2018_threshold, gravity;
  int32_t source_method; /* 0 semi-implicit, 1 XQ2018 */
  int32_t well_balancing_bs2002; /* 0: eta_vertices stays at the vertex elevations (dhv == 0, src/operator_fluxes_ceed.c:687-695) */
} Bs2002Params;

const char *bs2002_oracle_kind(void) { return "reference-ceed-qfunctions"; }

static void sort3(double *a, double *b, double *c) { /* src/operator_fluxes_ceed.c sort3: ascending */
  double t;
  if (*a > *b) { t = *a; *a = *b; *b = t; }
  if (*b > *c) { t = *b; *b = *c; *c = t; }
  if (*a > *b) { t = *a; *a = *b; *b = t; }
}

/* CreateCeedEtaVerticesOperator: eta_vertices[v] = sum over the cells c around v of SWEEtaVertex(z1<=z2<=z3 of c, q[c]) / n_cells(v),
 * connections visited vertex by vertex, the cells of a vertex in ascending cell id */
static void eta_vertices(const Bs2002Mesh *m, struct SWEContext_ *ctx, const double *u, const int32_t *vc_off, const int32_t *vc_ids, double *eta) {
  for (int32_t v = 0; v < m->num_vertices; ++v) {
    eta[v] = 0.0;
    const int32_t nc = vc_off[v + 1] - vc_off[v];
    for (int32_t k = vc_off[v]; k < vc_off[v + 1]; ++k) {
      const int32_t c  = vc_ids[k];
      const int32_t o  = m->cell_vertex_offsets[c];
      double        z1 = m->vertex_z[m->cell_vertex_ids[o]], z2 = m->vertex_z[m->cell_vertex_ids[o + 1]], z3 = m->vertex_z[m->cell_vertex_ids[o + 2]];
      sort3(&z1, &z2, &z3);
      /* the Q-function sees in[0] = geom as [4][Q], in[1] = q as [3][Q]; with Q = 1 a [n][1] array is n consecutive scalars */
      CeedScalar        gflat[4] = {z1, z2, z3, 1.0 / nc}, qflat[3] = {u[3 * c], u[3 * c + 1], u[3 * c + 2]}, out0;
      const CeedScalar *in[2]    = {gflat, qflat};
      CeedScalar       *out[1]   = {&out0};
      SWEEtaVertex(ctx, 1, in, out);
      eta[v] += out0;
    }
  }
}

/* One right-hand-side evaluation: F = flux divergence + sources, u and F AoS [num_cells][3].  eta (length num_vertices) returns the
 * vertex water-surface elevations that were used; courant returns the maximum Courant number. */
int bs2002_oracle_rhs(const Bs2002Mesh *m, const Bs2002Params *p, double dt, const double *u, const double *mannings, const double *ext_src, double *F,
                      double *eta, double *courant) {
  struct SWEContext_ ctx = {.dtime = dt, .tiny_h = p->tiny_h, .h_anuga_regular = p->h_anuga_regular, .gravity = p->gravity,
                            .xq2018_threshold = p->xq2018_threshold};
  const int32_t nc = m->num_cells, nv = m->num_vertices;
  /* vertex -> cells, ascending cell id */
  int32_t *vc_off = calloc((size_t)nv + 1, sizeof(int32_t)), *vc_ids, *fill;
  for (int32_t c = 0; c < nc; ++c)
    for (int32_t k = m->cell_vertex_offsets[c]; k < m->cell_vertex_offsets[c + 1]; ++k) vc_off[m->cell_vertex_ids[k] + 1]++;
  for (int32_t v = 0; v < nv; ++v) vc_off[v + 1] += vc_off[v];
  vc_ids = malloc(sizeof(int32_t) * (size_t)(vc_off[nv] > 0 ? vc_off[nv] : 1));
  fill   = malloc(sizeof(int32_t) * (size_t)(nv > 0 ? nv : 1));
  memcpy(fill, vc_off, sizeof(int32_t) * (size_t)nv);
  for (int32_t c = 0; c < nc; ++c)
    for (int32_t k = m->cell_vertex_offsets[c]; k < m->cell_vertex_offsets[c + 1]; ++k) vc_ids[fill[m->cell_vertex_ids[k]]++] = c;
  if (p->well_balancing_bs2002) eta_vertices(m, &ctx, u, vc_off, vc_ids, eta);
  else memcpy(eta, m->vertex_z, sizeof(double) * (size_t)nv); /* CreateCeedEtaVerticesVector initialises it with the vertex z: dhv == 0 */
  free(vc_off); free(vc_ids); free(fill);

  double cmax = 0.0;
  memset(F, 0, sizeof(double) * 3 * (size_t)nc);
  /* interior flux operator: Q-function per edge, then the transposes of the "cell_left" and "cell_right" restrictions, one after the other */
  const int32_t nie = m->num_internal_edges;
  double *cl = malloc(sizeof(double) * 3 * (size_t)(nie > 0 ? nie : 1)), *cr = malloc(sizeof(double) * 3 * (size_t)(nie > 0 ? nie : 1));
  for (int32_t i = 0; i < nie; ++i) {
    const int32_t e = m->internal_edge_ids[i], l = m->edge_cell_ids[2 * e], r = m->edge_cell_ids[2 * e + 1];
    const int32_t vb = m->edge_vertex_ids[2 * e], ve = m->edge_vertex_ids[2 * e + 1];
    CeedScalar geom[6] = {m->edge_sn[e], m->edge_cn[e], -m->edge_lengths[e] / m->cell_areas[l], m->edge_lengths[e] / m->cell_areas[r], m->vertex_z[vb],
                          m->vertex_z[ve]};
    CeedScalar ql[3] = {u[3 * l], u[3 * l + 1], u[3 * l + 2]}, qr[3] = {u[3 * r], u[3 * r + 1], u[3 * r + 2]}, eb = eta[vb], ee = eta[ve];
    CeedScalar flux[3], cn[2];
    const CeedScalar *in[5]  = {geom, ql, qr, &eb, &ee};
    CeedScalar       *out[4] = {cl + 3 * i, cr + 3 * i, flux, cn};
    if (SWEFlux_Roe(&ctx, 1, in, out)) return 1;
    if (fabs(cn[0]) > cmax) cmax = fabs(cn[0]);
    if (fabs(cn[1]) > cmax) cmax = fabs(cn[1]);
  }
  for (int32_t i = 0; i < nie; ++i) {
    const int32_t l = m->edge_cell_ids[2 * m->internal_edge_ids[i]];
    for (int k = 0; k < 3; ++k) F[3 * l + k] += cl[3 * i + k];
  }
  for (int32_t i = 0; i < nie; ++i) {
    const int32_t r = m->edge_cell_ids[2 * m->internal_edge_ids[i] + 1];
    for (int k = 0; k < 3; ++k) F[3 * r + k] += cr[3 * i + k];
  }
  free(cl); free(cr);
  /* boundary flux operators */
  for (int32_t j = 0; j < m->num_boundary_edges; ++j) {
    const int32_t e = m->boundary_edge_ids[j], l = m->edge_cell_ids[2 * e];
    const int32_t vb = m->edge_vertex_ids[2 * e], ve = m->edge_vertex_ids[2 * e + 1];
    CeedScalar geom[5] = {m->edge_sn[e], m->edge_cn[e], -m->edge_lengths[e] / m->cell_areas[l], m->vertex_z[vb], m->vertex_z[ve]};
    CeedScalar ql[3] = {u[3 * l], u[3 * l + 1], u[3 * l + 2]}, eb = eta[vb], ee = eta[ve], cell[3], flux[3], cn[1];
    CeedScalar *out[3] = {cell, flux, cn};
    int         ierr;
    if (m->boundary_edge_type[j] == 0) {
      CeedScalar qr[3] = {m->boundary_values[3 * j], m->boundary_values[3 * j + 1], m->boundary_values[3 * j + 2]};
      const CeedScalar *in[5] = {geom, ql, &eb, &ee, qr};
      ierr = SWEBoundaryFlux_Dirichlet_Roe(&ctx, 1, in, out);
    } else if (m->boundary_edge_type[j] == 2) {
      const CeedScalar *in[4] = {geom, ql, &eb, &ee};
      ierr = SWEBoundaryFlux_Reflecting_Roe(&ctx, 1, in, out);
    } else {
      const CeedScalar *in[4] = {geom, ql, &eb, &ee};
      ierr = SWEBoundaryFlux_Outflow_Roe(&ctx, 1, in, out);
    }
    if (ierr) return 1;
    for (int k = 0; k < 3; ++k) F[3 * l + k] += cell[k];
    if (fabs(cn[0]) > cmax) cmax = fabs(cn[0]);
  }
  /* source operator: riemannf = the flux divergence accumulated above (src/operator_sources_ceed.c), output = the complete RHS */
  for (int32_t c = 0; c < nc; ++c) {
    CeedScalar geom[2] = {m->cell_dz_dx[c], m->cell_dz_dy[c]}, src[3] = {ext_src[3 * c], ext_src[3 * c + 1], ext_src[3 * c + 2]}, mat[1] = {mannings[c]};
    CeedScalar rf[3] = {F[3 * c], F[3 * c + 1], F[3 * c + 2]}, q[3] = {u[3 * c], u[3 * c + 1], u[3 * c + 2]}, s[3], prim[3], inst[3];
    const CeedScalar *in[5]  = {geom, src, mat, rf, q};
    CeedScalar       *out[3] = {s, prim, inst};
    if (p->source_method == 1) SWESourcesWithImplicitBedFrictionXQ2018(&ctx, 1, in, out);
    else SWESourcesWithSemiImplicitBedFriction(&ctx, 1, in, out);
    for (int k = 0; k < 3; ++k) F[3 * c + k] = s[k];
  }
  if (courant) *courant = cmax;
  return 0;
}

/* nsteps forward-Euler steps (TSEULER: u + dt F, unfused; third-party arithmetic) */
int bs2002_oracle_step_euler(const Bs2002Mesh *m, const Bs2002Params *p, double dt, int32_t nsteps, double *u, const double *mannings, const double *ext_src,
                             double *eta, double *courant) {
  double *F = malloc(sizeof(double) * 3 * (size_t)m->num_cells);
  for (int32_t s = 0; s < nsteps; ++s) {
    if (bs2002_oracle_rhs(m, p, dt, u, mannings, ext_src, F, eta, courant)) { free(F); return 1; }
    for (int64_t i = 0; i < 3 * (int64_t)m->num_cells; ++i) u[i] = u[i] + dt * F[i];
  }
  free(F);
  return 0;
}
