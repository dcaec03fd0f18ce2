/* ref_driver.c — drives the REFERENCE's own PETSc-backend SWE operator.  TEST INFRASTRUCTURE ONLY.
 *
 * Linked with the reference sources compiled in place from /root/reference (see oracle/Makefile):
 *   src/swe/swe_petsc.c (+ swe_roe_flux_petsc.h), src/petsc_operator.c,
 *   src/operator_fluxes_petsc.c, src/operator_sources_petsc.c
 * against oracle/ref_shim/ (serial stand-in for PETSc/libCEED/MPI types).  This file only
 *   - fills RDyMesh / RDyConfig / RDyBoundary / RDyCondition from an RdybMeshView,
 *   - calls the reference factories CreatePetscFlux[HR]Operator / CreatePetscSource[HR]Operator
 *     exactly as CreateOperatorSubOperators does (src/operator.c:195-216),
 *   - restates the four statements of ApplyPetscOperator (src/operator.c:656-672), the serial part of
 *     OperatorRHSFunction (src/rdysetup.c:1127-1139) and PETSc's TSEULER / TSRK4 stage updates
 *     (third-party, not under /root/reference: "parity unpinned" at that boundary).
 * No reference source is copied into this repository.
 */
#include <private/rdycoreimpl.h>
#include <private/rdyoperatorimpl.h>
#include <private/rdysweimpl.h>
#include <private/rdytracerimpl.h>

#include "swe_oracle.h"

/* ---- symbols the four reference files reference but the SWE first-order path never reaches ---- */
#define UNREACHED(name) \
  { fprintf(stderr, "ref_driver: %s is outside the oracle's scope\n", name); return PETSC_ERR_SUP_SHIM; }
#define PETSC_ERR_SUP_SHIM 56
PetscErrorCode CreatePetscTracerInteriorFluxOperator(RDyMesh *m, const RDyConfig c, OperatorDiagnostics *d, PetscOperator *o) UNREACHED("tracers")
PetscErrorCode CreatePetscTracerInteriorFluxHROperator(RDyMesh *m, const RDyConfig c, OperatorDiagnostics *d, PetscOperator *o) UNREACHED("tracers")
PetscErrorCode CreatePetscTracerBoundaryFluxOperator(RDyMesh *m, const RDyConfig c, RDyBoundary b, RDyCondition bc, Vec v, Vec f,
                                                     OperatorDiagnostics *d, PetscOperator *o) UNREACHED("tracers")
PetscErrorCode CreatePetscTracerSourceOperator(RDyMesh *m, const RDyConfig c, Vec e, Vec p, PetscOperator *o) UNREACHED("tracers")
PetscErrorCode CreatePetscTracerSourceHROperator(RDyMesh *m, const RDyConfig c, Vec e, Vec p, PetscOperator *o) UNREACHED("tracers")

/* ---- the four second-order helpers of src/operator_fluxes_ceed.c (a libCEED translation unit that cannot be built here),
 * restated in muscl_helpers.h and bound to the reference's signatures so that the reference's own ApplyInteriorFlux2R
 * (src/swe/swe_petsc.c:98-213) runs on top of them ---- */
#include "muscl_helpers.h"
static void mesh_xy(RDyMesh *m, double **xc, double **yc, double **xm, double **ym) {
  *xc = malloc(sizeof(double) * (size_t)(m->num_cells + 1)); *yc = malloc(sizeof(double) * (size_t)(m->num_cells + 1));
  for (PetscInt c = 0; c < m->num_cells; ++c) { (*xc)[c] = m->cells.centroids[c].X[0]; (*yc)[c] = m->cells.centroids[c].X[1]; }
  if (!xm) return;
  *xm = malloc(sizeof(double) * (size_t)(m->num_edges + 1)); *ym = malloc(sizeof(double) * (size_t)(m->num_edges + 1));
  for (PetscInt e = 0; e < m->num_edges; ++e) { /* edge midpoint from its two vertices, :1162-1166 */
    PetscInt v0 = m->edges.vertex_ids[2 * e], v1 = m->edges.vertex_ids[2 * e + 1];
    (*xm)[e] = 0.5 * (m->vertices.points[v0].X[0] + m->vertices.points[v1].X[0]);
    (*ym)[e] = 0.5 * (m->vertices.points[v0].X[1] + m->vertices.points[v1].X[1]);
  }
}
PetscErrorCode PrecomputeLSGradCoeffs(MPI_Comm comm, RDyMesh *m, PetscReal *c) {
  (void)comm;
  double *xc, *yc;
  mesh_xy(m, &xc, &yc, NULL, NULL);
  int ierr = muscl_ls_coeffs(m->num_cells, m->num_internal_edges, m->edges.internal_edge_ids, m->edges.cell_ids, xc, yc, c);
  free(xc); free(yc);
  return ierr;
}
PetscErrorCode ComputeLeastSquaresGradients(RDyMesh *m, const PetscReal *c, const PetscScalar *u, PetscScalar *a, PetscScalar *b,
                                            PetscScalar *d) {
  muscl_gradients(m->num_cells, m->num_internal_edges, m->edges.internal_edge_ids, m->edges.cell_ids, c, u, a, b, d);
  return 0;
}
/* serial: the owned -> ghost round trip leaves the gradients unchanged (:1049-1051) */
PetscErrorCode CommunicateCellGradients(DM dm, RDyMesh *m, PetscScalar *a, PetscScalar *b, PetscScalar *c) {
  (void)dm; (void)m; (void)a; (void)b; (void)c;
  return 0;
}
PetscErrorCode ReconstructFaceValues(RDyMesh *m, const PetscScalar *u, const PetscScalar *a, const PetscScalar *b, const PetscScalar *c,
                                     RDyLimiterType l, CeedScalar *q) {
  double *xc, *yc, *xm, *ym;
  mesh_xy(m, &xc, &yc, &xm, &ym);
  muscl_faces(m->num_internal_edges, m->edges.internal_edge_ids, m->edges.cell_ids, (const int32_t *)m->edges.is_owned, xc, yc, xm, ym, u, a, b,
              c, (int)l, q);
  free(xc); free(yc); free(xm); free(ym);
  return 0;
}

struct SweOracle_ {
  RDyConfig        *config;
  RDyMesh           mesh;
  PetscInt          nb;
  RDyBoundary      *boundaries;
  RDyCondition     *conditions;
  RDyFlowCondition *flows;
  Operator          op; /* the reference's Operator struct; only the members ApplyPetscOperator touches are set */
  Vec               u_global, u_local, rhs;
  Vec               k[4], ystage; /* RK4 work vectors */
  struct _shim_DM   dm;           /* serial DM behind u_local (second-order path) */
};

static Vec vec_new(PetscInt n, PetscInt bs) {
  Vec v = calloc(1, sizeof(*v));
  v->a  = calloc((size_t)(n > 0 ? n : 1), sizeof(PetscScalar));
  v->n  = n;
  v->bs = bs;
  return v;
}
static void vec_free(Vec v) {
  if (v) { free(v->a); free(v); }
}

#ifdef RDYB_ORACLE_USE_B200_FACTORIES
/* the SAME driver with the four factory calls of CreateOperatorSubOperators switched to the B200 adapter
 * (rdycore_b200/csrc/rdycore_adapter.c, the reference's signatures): oracle/_ref/libswe_adapter_ref.so, kind "adapter".  Everything
 * else - the Operator struct and its Vecs, PetscOperatorApply / SetField of the reference's src/petsc_operator.c, the four
 * statements of ApplyPetscOperator - is unchanged, so a test can compare the two libraries array by array. */
PetscErrorCode CreateB200FluxOperator(RDyConfig *, RDyMesh *, MPI_Comm, PetscInt, RDyBoundary *, RDyCondition *, Vec *, Vec *, Vec *, OperatorDiagnostics *,
                                      PetscOperator *);
PetscErrorCode CreateB200FluxHROperator(RDyConfig *, RDyMesh *, PetscInt, RDyBoundary *, RDyCondition *, Vec *, Vec *, Vec *, OperatorDiagnostics *,
                                        PetscOperator *);
PetscErrorCode CreateB200SourceOperator(RDyConfig *, RDyMesh *, Vec, Vec, PetscOperator *);
PetscErrorCode CreateB200SourceHROperator(RDyConfig *, RDyMesh *, Vec, Vec, PetscOperator *);
#define CreatePetscFluxOperator CreateB200FluxOperator
#define CreatePetscFluxHROperator CreateB200FluxHROperator
#define CreatePetscSourceOperator CreateB200SourceOperator
#define CreatePetscSourceHROperator CreateB200SourceHROperator
const char *swe_oracle_kind(void) { return "adapter"; }
#else
const char *swe_oracle_kind(void) { return "reference"; }
#endif

int swe_oracle_create(const RdybConfig *cfg, const RdybMeshView *mv, int32_t nb, const RdybBoundaryView *bv, SweOracle *out) {
  SweOracle o = calloc(1, sizeof(*o));
  *out        = o;
  o->config   = calloc(1, sizeof(RDyConfig));
  RDyConfig *c = o->config;
  c->physics.flow.mode                    = FLOW_SWE;
  c->physics.flow.tiny_h                  = cfg->tiny_h;
  c->physics.flow.h_anuga_regular         = cfg->h_anuga_regular;
  c->physics.flow.source.method           = (RDyFlowSourceMethod)cfg->source_method;
  c->physics.flow.source.xq2018_threshold = cfg->xq2018_threshold;
  c->physics.flow.well_balancing          = (RDyWellBalanceMethod)cfg->well_balancing;
  c->numerics.riemann                     = RIEMANN_ROE;
  c->numerics.temporal                    = TEMPORAL_EULER;
  c->numerics.second_order                = cfg->reconstruction != RDYB_RECON_FIRST_ORDER;
  c->numerics.limiter = cfg->reconstruction == RDYB_RECON_MUSCL_NONE ? LIMITER_NONE : cfg->reconstruction == RDYB_RECON_MUSCL_VANLEER ? LIMITER_VANLEER : LIMITER_MINMOD;

  /* RDyMesh from the view: only the members the PETSc SWE operators read */
  RDyMesh *m = &o->mesh;
  m->num_cells = mv->num_cells; m->num_owned_cells = mv->num_owned_cells; m->num_edges = mv->num_edges;
  m->num_internal_edges = mv->num_internal_edges;
  m->cells.is_owned       = (PetscBool *)mv->cell_is_owned;
  m->cells.local_to_owned = (PetscInt *)mv->cell_local_to_owned;
  m->cells.areas          = (PetscReal *)mv->cell_areas;
  m->cells.dz_dx          = (PetscReal *)mv->cell_dz_dx;
  m->cells.dz_dy          = (PetscReal *)mv->cell_dz_dy;
  m->cells.global_ids     = calloc((size_t)mv->num_cells, sizeof(PetscInt));
  for (PetscInt i = 0; i < mv->num_cells; ++i) m->cells.global_ids[i] = mv->cell_global_ids ? (PetscInt)mv->cell_global_ids[i] : i;
  m->edges.cell_ids          = (PetscInt *)mv->edge_cell_ids;
  m->edges.lengths           = (PetscReal *)mv->edge_lengths;
  m->edges.cn                = (PetscReal *)mv->edge_cn;
  m->edges.sn                = (PetscReal *)mv->edge_sn;
  m->edges.internal_edge_ids = (PetscInt *)mv->internal_edge_ids;
  m->edges.global_ids        = calloc((size_t)(mv->num_edges > 0 ? mv->num_edges : 1), sizeof(PetscInt));
  for (PetscInt e = 0; e < mv->num_edges; ++e) m->edges.global_ids[e] = mv->edge_global_ids ? (PetscInt)mv->edge_global_ids[e] : e;
  if (cfg->well_balancing == RDYB_WELL_BALANCING_HR) {
    /* take the reference's "cell_elevation file" branch (swe_petsc.c:1213-1215): zc = centroid z */
    if (!mv->cell_zc) { fprintf(stderr, "ref_driver: HR needs cell_zc\n"); return PETSC_ERR_USER; }
    strcpy(c->grid.cell_elevation.file, "from-mesh-view");
    m->cells.centroids = calloc((size_t)mv->num_cells, sizeof(RDyPoint));
    for (PetscInt i = 0; i < mv->num_cells; ++i) m->cells.centroids[i].X[2] = mv->cell_zc[i];
  }
  if (c->numerics.second_order && cfg->well_balancing == RDYB_WELL_BALANCING_HR) {
    /* CreateOperator's check, src/operator.c:388-389 (operator.c itself is not part of this build) */
    fprintf(stderr, "-second_order cannot be used with well_balancing = HR simultaneously (not yet implemented)\n");
    return PETSC_ERR_USER;
  }
  if (c->numerics.second_order) {
    /* what ApplyInteriorFlux2R and its helpers read: centroids, edge vertices (two synthetic vertices per edge, both AT the
     * view's midpoint, so that 0.5*(v0+v1) reproduces it exactly), edge ownership (serial: every edge is owned) */
    if (!mv->cell_xc || !mv->cell_yc || !mv->edge_xm || !mv->edge_ym) { fprintf(stderr, "ref_driver: second order needs centroids and edge midpoints\n"); return PETSC_ERR_USER; }
    if (!m->cells.centroids) m->cells.centroids = calloc((size_t)mv->num_cells, sizeof(RDyPoint));
    for (PetscInt i = 0; i < mv->num_cells; ++i) { m->cells.centroids[i].X[0] = mv->cell_xc[i]; m->cells.centroids[i].X[1] = mv->cell_yc[i]; }
    size_t ne = (size_t)(mv->num_edges > 0 ? mv->num_edges : 1);
    m->num_vertices     = 2 * mv->num_edges;
    m->vertices.points  = calloc(2 * ne, sizeof(RDyPoint));
    m->edges.vertex_ids = calloc(2 * ne, sizeof(PetscInt));
    m->edges.is_owned   = calloc(ne, sizeof(PetscBool));
    for (PetscInt e = 0; e < mv->num_edges; ++e) {
      for (int k = 0; k < 2; ++k) {
        m->edges.vertex_ids[2 * e + k] = 2 * e + k;
        m->vertices.points[2 * e + k].X[0] = mv->edge_xm[e]; m->vertices.points[2 * e + k].X[1] = mv->edge_ym[e];
      }
      m->edges.is_owned[e] = PETSC_TRUE;
    }
    m->num_owned_internal_edges = mv->num_internal_edges;
  }

  o->nb         = nb;
  o->boundaries = calloc((size_t)(nb > 0 ? nb : 1), sizeof(RDyBoundary));
  o->conditions = calloc((size_t)(nb > 0 ? nb : 1), sizeof(RDyCondition));
  o->flows      = calloc((size_t)(nb > 0 ? nb : 1), sizeof(RDyFlowCondition));
  for (PetscInt b = 0; b < nb; ++b) {
    o->boundaries[b].id = bv[b].id; o->boundaries[b].index = bv[b].index; o->boundaries[b].num_edges = bv[b].num_edges;
    o->boundaries[b].edge_ids = (PetscInt *)bv[b].edge_ids;
    o->flows[b].type          = (RDyConditionType)bv[b].bc_type;
    o->conditions[b].flow     = &o->flows[b];
  }

  /* SetOperatorBoundaries / SetOperatorRegions (src/operator.c:78-133): the sequential Vecs */
  Operator *op            = &o->op;
  op->config              = c;
  op->num_components      = 3;
  op->mesh                = m;
  op->num_boundaries      = nb;
  op->boundaries          = o->boundaries;
  op->boundary_conditions = o->conditions;
  op->petsc.boundary_values       = calloc((size_t)(nb > 0 ? nb : 1), sizeof(Vec));
  op->petsc.boundary_fluxes       = calloc((size_t)(nb > 0 ? nb : 1), sizeof(Vec));
  op->petsc.boundary_fluxes_accum = calloc((size_t)(nb > 0 ? nb : 1), sizeof(Vec));
  for (PetscInt b = 0; b < nb; ++b) {
    op->petsc.boundary_values[b]       = vec_new(3 * bv[b].num_edges, 3);
    op->petsc.boundary_fluxes[b]       = vec_new(3 * bv[b].num_edges, 3);
    op->petsc.boundary_fluxes_accum[b] = vec_new(3 * bv[b].num_edges, 3);
  }
  PetscInt N                    = mv->num_owned_cells;
  op->petsc.external_sources    = vec_new(3 * N, 3);
  op->petsc.material_properties = vec_new(N, NUM_MATERIAL_PROPERTIES);

  /* CreateOperatorSubOperators, PETSc branch (src/operator.c:195-216) */
  if (cfg->well_balancing == RDYB_WELL_BALANCING_HR) {
    PetscCall(CreatePetscFluxHROperator(c, m, nb, o->boundaries, o->conditions, op->petsc.boundary_values, op->petsc.boundary_fluxes,
                                        op->petsc.boundary_fluxes_accum, &op->diagnostics, &op->petsc.flux));
    PetscCall(CreatePetscSourceHROperator(c, m, op->petsc.external_sources, op->petsc.material_properties, &op->petsc.source));
  } else {
    PetscCall(CreatePetscFluxOperator(c, m, PETSC_COMM_WORLD, nb, o->boundaries, o->conditions, op->petsc.boundary_values,
                                      op->petsc.boundary_fluxes, op->petsc.boundary_fluxes_accum, &op->diagnostics, &op->petsc.flux));
    PetscCall(CreatePetscSourceOperator(c, m, op->petsc.external_sources, op->petsc.material_properties, &op->petsc.source));
  }
  /* AddOperatorFluxDivergence / PrimitiveVariables / SourceVariables (src/operator.c:224-335) */
  op->flux_divergence     = vec_new(3 * N, 3);
  op->primitive_variables = vec_new(3 * N, 3);
  op->src_inst            = vec_new(3 * N, 3);
  PetscCall(PetscOperatorSetField(op->petsc.source, "riemannf", op->flux_divergence));
  PetscCall(PetscOperatorSetField(op->petsc.source, "primitive_variables", op->primitive_variables));

  o->u_global = vec_new(3 * N, 3);
  o->rhs      = vec_new(3 * N, 3);
  o->u_local  = vec_new(3 * mv->num_cells, 3);
  o->dm = (struct _shim_DM){.n_local = mv->num_cells, .bs = 3, .is_owned = m->cells.is_owned, .local_to_owned = m->cells.local_to_owned};
  o->u_local->dm = &o->dm;
  for (int i = 0; i < 4; ++i) o->k[i] = vec_new(3 * N, 3);
  o->ystage = vec_new(3 * N, 3);
  return 0;
}

void swe_oracle_destroy(SweOracle o) {
  if (!o) return;
  Operator *op = &o->op;
  if (op->petsc.flux) PetscOperatorDestroy(&op->petsc.flux);
  if (op->petsc.source) PetscOperatorDestroy(&op->petsc.source);
  for (PetscInt b = 0; b < o->nb; ++b) {
    vec_free(op->petsc.boundary_values[b]); vec_free(op->petsc.boundary_fluxes[b]); vec_free(op->petsc.boundary_fluxes_accum[b]);
  }
  free(op->petsc.boundary_values); free(op->petsc.boundary_fluxes); free(op->petsc.boundary_fluxes_accum);
  vec_free(op->petsc.external_sources); vec_free(op->petsc.material_properties);
  vec_free(op->flux_divergence); vec_free(op->primitive_variables); vec_free(op->src_inst);
  vec_free(o->u_global); vec_free(o->rhs); vec_free(o->u_local); vec_free(o->ystage);
  for (int i = 0; i < 4; ++i) vec_free(o->k[i]);
  free(o->mesh.cells.global_ids); free(o->mesh.edges.global_ids); free(o->mesh.cells.centroids);
  free(o->mesh.vertices.points); free(o->mesh.edges.vertex_ids); free(o->mesh.edges.is_owned);
  free(o->boundaries); free(o->conditions); free(o->flows); free(o->config);
  free(o);
}

double *swe_oracle_array(SweOracle o, const char *name, int32_t b, int64_t *length) {
  Vec       v  = NULL;
  Operator *op = &o->op;
  if (!strcmp(name, "u_global")) v = o->u_global;
  else if (!strcmp(name, "u_local")) v = o->u_local;
  else if (!strcmp(name, "rhs")) v = o->rhs;
  else if (!strcmp(name, "external_sources")) v = op->petsc.external_sources;
  else if (!strcmp(name, "material_properties")) v = op->petsc.material_properties;
  else if (!strcmp(name, "flux_divergence")) v = op->flux_divergence;
  else if (!strcmp(name, "primitive_variables")) v = op->primitive_variables;
  else if (!strcmp(name, "src_inst")) v = op->src_inst;
  else if (b >= 0 && b < o->nb) {
    if (!strcmp(name, "boundary_values")) v = op->petsc.boundary_values[b];
    else if (!strcmp(name, "boundary_fluxes")) v = op->petsc.boundary_fluxes[b];
    else if (!strcmp(name, "boundary_fluxes_accum")) v = op->petsc.boundary_fluxes_accum[b];
  }
  if (!v) return NULL;
  if (length) *length = v->n;
  return v->a;
}

/* F(U): OperatorRHSFunction body, src/rdysetup.c:1127-1139, with ApplyPetscOperator src/operator.c:656-672 */
static PetscErrorCode rhs_function(SweOracle o, PetscReal dt, Vec U, Vec F) {
  Operator *op = &o->op;
  RDyMesh  *m  = &o->mesh;
  PetscCall(VecZeroEntries(F));
  /* DMGlobalToLocal(INSERT), serial part: owned entries; ghosts are filled by the test harness */
  for (PetscInt c = 0; c < m->num_cells; ++c) {
    if (m->cells.is_owned[c]) {
      PetscInt oc = m->cells.local_to_owned[c];
      for (int d = 0; d < 3; ++d) o->u_local->a[3 * c + d] = U->a[3 * oc + d];
    }
  }
  /* ResetOperatorDiagnostics, src/operator.c:757-768 */
  op->diagnostics.updated        = PETSC_FALSE;
  op->diagnostics.courant_number = (CourantNumberDiagnostics){.max_courant_num = 0.0, .global_edge_id = -1, .global_cell_id = -1};
  PetscCall(PetscOperatorApply(op->petsc.flux, dt, o->u_local, F));
  PetscCall(VecCopy(F, op->flux_divergence));
  PetscCall(PetscOperatorApply(op->petsc.source, dt, o->u_local, F));
  PetscCall(VecCopy(op->petsc.external_sources, op->src_inst));
  return 0;
}

int swe_oracle_rhs(SweOracle o, double dt) { return rhs_function(o, dt, o->u_global, o->rhs); }

int swe_oracle_euler_update(SweOracle o, double dt) {
  /* PETSc TSStep_Euler: VecAYPX(update, dt, solution) i.e. update = solution + dt*update; VecCopy(update, solution) */
  for (PetscInt i = 0; i < o->u_global->n; ++i) o->u_global->a[i] = o->u_global->a[i] + dt * o->rhs->a[i];
  return 0;
}

int swe_oracle_step_euler(SweOracle o, double dt, int32_t nsteps) {
  for (int32_t s = 0; s < nsteps; ++s) {
    PetscCall(swe_oracle_rhs(o, dt));
    PetscCall(swe_oracle_euler_update(o, dt));
  }
  return 0;
}

int swe_oracle_step_rk4(SweOracle o, double dt, int32_t nsteps) {
  /* classical RK4 tableau as PETSc TSRK4: Y_i = U + dt*sum_j A_ij K_j (VecMAXPY order), U += dt*sum_j b_j K_j */
  static const double A[4][3] = {{0, 0, 0}, {0.5, 0, 0}, {0, 0.5, 0}, {0, 0, 1.0}};
  static const double bw[4]   = {1.0 / 6.0, 1.0 / 3.0, 1.0 / 3.0, 1.0 / 6.0};
  PetscInt n = o->u_global->n;
  for (int32_t s = 0; s < nsteps; ++s) {
    for (int i = 0; i < 4; ++i) {
      PetscCall(VecCopy(o->u_global, o->ystage));
      for (int j = 0; j < i; ++j) {
        double w = dt * A[i][j];
        if (w != 0.0) for (PetscInt q = 0; q < n; ++q) o->ystage->a[q] += w * o->k[j]->a[q];
      }
      PetscCall(rhs_function(o, dt, o->ystage, o->k[i]));
    }
    for (int j = 0; j < 4; ++j) {
      double w = dt * bw[j];
      for (PetscInt q = 0; q < n; ++q) o->u_global->a[q] += w * o->k[j]->a[q];
    }
  }
  return 0;
}

int swe_oracle_courant(SweOracle o, RdybCourantDiagnostics *d) {
  d->max_courant_num = o->op.diagnostics.courant_number.max_courant_num;
  d->global_edge_id  = o->op.diagnostics.courant_number.global_edge_id;
  d->global_cell_id  = o->op.diagnostics.courant_number.global_cell_id;
  return 0;
}
