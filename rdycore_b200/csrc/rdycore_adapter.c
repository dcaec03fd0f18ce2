/* rdycore_adapter.c — the B200 SWE path behind RDycore's sub-operator factories, WITH THE REFERENCE'S OWN SIGNATURES.
 *
 * This translation unit is compiled INSIDE an RDycore build (or against RDycore's headers): it includes
 * <private/rdycoreimpl.h> / <private/rdyoperatorimpl.h> and speaks RDyConfig, RDyMesh, RDyBoundary, RDyCondition, Vec,
 * PetscOperator and OperatorDiagnostics natively.  It is NOT part of librdyb200.so (which must build without RDycore); it links
 * against it.  Exports, argument for argument the factories of include/private/rdyoperatorimpl.h:234-239:
 *
 *   CreateB200FluxOperator     <->  CreatePetscFluxOperator    (RDyConfig*, RDyMesh*, MPI_Comm, PetscInt, RDyBoundary*, RDyCondition*,
 *                                                               Vec*, Vec*, Vec*, OperatorDiagnostics*, PetscOperator*)
 *   CreateB200FluxHROperator   <->  CreatePetscFluxHROperator  (same without the communicator)
 *   CreateB200SourceOperator   <->  CreatePetscSourceOperator  (RDyConfig*, RDyMesh*, Vec, Vec, PetscOperator*)
 *   CreateB200SourceHROperator <->  CreatePetscSourceHROperator
 *
 * so the patch to CreateOperatorSubOperators (src/operator.c:195-216) is a rename of four calls under `if (b200.enable)`.
 * With -DRDYB_ADAPTER_REPLACE_PETSC the reference's names themselves are defined here (drop operator_fluxes_petsc.c /
 * operator_sources_petsc.c from the build instead of patching operator.c).
 *
 * What the factories do: translate the borrowed RDycore objects into the views of include/rdyb200.h (FillMeshView below: pointer
 * assignments plus the three arrays RDyMesh stores in another shape - centroids, 64-bit ids, the HR bed elevation
 * of src/swe/swe_petsc.c:1208-1224), create ONE device operator per RDyMesh, and wrap it in composite PetscOperators made by the
 * reference's own PetscOperatorCreate / CreateComposite / CompositeAddSub (src/petsc_operator.c), whose apply callbacks are the
 * shared sub-operators of b200_suboperators.h.  The flux and source operator of one Operator find each other through the RDyMesh
 * pointer they are both created with (CreateOperatorSubOperators passes op->mesh to both).
 *
 * Tested in this repository against the reference's headers behind oracle/ref_shim (oracle/Makefile target `adapter`,
 * tests/test_gpu_rdycore_adapter.py): the four statements of ApplyPetscOperator (src/operator.c:656-672) run once on the
 * reference's sub-operators and once on these, on the same Operator Vecs.
 */
#include <private/rdycoreimpl.h>
#include <private/rdyoperatorimpl.h>

#include "rdyb200.h"
#include "rdyb_error.h"

/* ---- bind b200_suboperators.h to the reference's types ---- */
#define B200_POP PetscOperator
#define B200_POPFIELDS PetscOperatorFields
#define B200_POP_CREATE PetscOperatorCreate
#define B200_POP_DESTROY PetscOperatorDestroy
#define B200_POP_CREATE_COMPOSITE PetscOperatorCreateComposite
#define B200_POP_ADD_SUB PetscOperatorCompositeAddSub
#define B200_POP_FIELDS_GET PetscOperatorFieldsGet
#define B200_DIAGS_T OperatorDiagnostics
/* running max; first-seen wins ties (src/swe/swe_petsc.c:289-296).  The ids are the view's global ids. */
#define B200_DIAGS_STORE(ptr, d)                                                        \
  do {                                                                                  \
    if ((d).max_courant_num > (ptr)->courant_number.max_courant_num) {                  \
      (ptr)->courant_number.max_courant_num = (d).max_courant_num;                      \
      (ptr)->courant_number.global_edge_id  = (PetscInt)(d).global_edge_id;             \
      (ptr)->courant_number.global_cell_id  = (PetscInt)(d).global_cell_id;             \
    }                                                                                   \
  } while (0)
static void adapter_unregister(void *shared);
#define B200_ON_RELEASE(s) adapter_unregister(s)

static int VecLocalSize64(Vec v, int64_t *n) {
  PetscInt       m;
  PetscErrorCode e = VecGetLocalSize(v, &m);
  *n               = m;
  return (int)e;
}
#ifdef RDYB_VEC_NO_STATE
/* a Vec implementation without an object state counter (the test shim): report a new state every time, i.e. re-upload the
 * operator data before every apply */
static int VecState64(Vec v, int64_t *s) {
  static int64_t counter = 0;
  (void)v;
  *s = ++counter;
  return 0;
}
#else
static int VecState64(Vec v, int64_t *s) {
  PetscObjectState t;
  PetscErrorCode   e = PetscObjectStateGet((PetscObject)v, &t);
  *s                 = (int64_t)t;
  return (int)e;
}
#endif
#include "b200_suboperators.h"

/* ---- one device operator per RDyMesh: the flux factory registers it, the source factory picks it up ---- */
#define ADAPTER_MAX_OPERATORS 64
static struct {
  const void *mesh;
  B200Shared *shared;
} g_registry[ADAPTER_MAX_OPERATORS];

static void adapter_unregister(void *shared) {
  for (int i = 0; i < ADAPTER_MAX_OPERATORS; ++i)
    if (g_registry[i].shared == shared) { g_registry[i].mesh = NULL; g_registry[i].shared = NULL; }
}
static int adapter_register(const void *mesh, B200Shared *s) {
  for (int i = 0; i < ADAPTER_MAX_OPERATORS; ++i)
    if (g_registry[i].mesh == mesh) { g_registry[i].shared = s; return RDYB_SUCCESS; } /* a new Operator on the same mesh replaces the entry */
  for (int i = 0; i < ADAPTER_MAX_OPERATORS; ++i)
    if (!g_registry[i].mesh) { g_registry[i].mesh = mesh; g_registry[i].shared = s; return RDYB_SUCCESS; }
  return rdyb_fail(RDYB_ERR_SUP, "more than %d live B200 operators", ADAPTER_MAX_OPERATORS);
}
static B200Shared *adapter_lookup(const void *mesh) {
  for (int i = 0; i < ADAPTER_MAX_OPERATORS; ++i)
    if (g_registry[i].mesh == mesh) return g_registry[i].shared;
  return NULL;
}

/* ---- RDycore objects -> views ---- */
typedef struct {
  RdybMeshView view;
  int64_t     *cell_gid, *edge_gid;
  double      *xc, *yc, *zc, *xm, *ym;
  int32_t     *is_owned;
} MeshViewStorage;

static void FreeMeshView(MeshViewStorage *st) {
  free(st->cell_gid); free(st->edge_gid); free(st->xc); free(st->yc); free(st->zc); free(st->xm); free(st->ym); free(st->is_owned);
  memset(st, 0, sizeof(*st));
}

/* The arrays the path consumes, borrowed from RDyMesh (include/private/rdymeshimpl.h:26-140) where the layout agrees and converted
 * where it does not.  hr: also the bed elevation of CreatePetscSWEInteriorFluxHROperator (src/swe/swe_petsc.c:1208-1224).
 * second_order: also centroids and edge midpoints as ReconstructFaceValues forms them (src/operator_fluxes_ceed.c:1162-1166). */
static int FillMeshView(const RDyConfig *config, RDyMesh *mesh, int hr, int second_order, MeshViewStorage *st) {
  memset(st, 0, sizeof(*st));
  RDYB_CHECK(sizeof(PetscInt) == sizeof(int32_t), RDYB_ERR_SUP, "the B200 path uses 32-bit local indices: configure PETSc without --with-64-bit-indices");
  RDYB_CHECK(sizeof(PetscReal) == sizeof(double), RDYB_ERR_SUP, "the B200 path is fp64: PetscReal must be double");
  RdybMeshView *v        = &st->view;
  const size_t  nc = (size_t)(mesh->num_cells > 0 ? mesh->num_cells : 1), ne = (size_t)(mesh->num_edges > 0 ? mesh->num_edges : 1);
  v->num_cells           = mesh->num_cells;
  v->num_owned_cells     = mesh->num_owned_cells;
  v->num_edges           = mesh->num_edges;
  v->num_internal_edges  = mesh->num_internal_edges;
  v->cell_local_to_owned = (const int32_t *)mesh->cells.local_to_owned;
  v->cell_areas          = mesh->cells.areas;
  v->cell_dz_dx          = mesh->cells.dz_dx;
  v->cell_dz_dy          = mesh->cells.dz_dy;
  v->edge_cell_ids       = (const int32_t *)mesh->edges.cell_ids;
  v->edge_lengths        = mesh->edges.lengths;
  v->edge_cn             = mesh->edges.cn;
  v->edge_sn             = mesh->edges.sn;
  v->internal_edge_ids   = (const int32_t *)mesh->edges.internal_edge_ids;
  /* PetscBool is an enum: copy instead of assuming its width */
  st->is_owned = malloc(sizeof(int32_t) * nc);
  st->cell_gid = malloc(sizeof(int64_t) * nc);
  st->edge_gid = malloc(sizeof(int64_t) * ne);
  RDYB_CHECK(st->is_owned && st->cell_gid && st->edge_gid, RDYB_ERR_MEM, "out of memory");
  for (PetscInt c = 0; c < mesh->num_cells; ++c) {
    st->is_owned[c] = mesh->cells.is_owned[c] ? 1 : 0;
    st->cell_gid[c] = mesh->cells.global_ids ? (int64_t)mesh->cells.global_ids[c] : (int64_t)c;
  }
  for (PetscInt e = 0; e < mesh->num_edges; ++e) st->edge_gid[e] = mesh->edges.global_ids ? (int64_t)mesh->edges.global_ids[e] : (int64_t)e;
  v->cell_is_owned   = st->is_owned;
  v->cell_global_ids = st->cell_gid;
  v->edge_global_ids = st->edge_gid;
  if (mesh->cells.centroids) { /* used to order the cells along a space-filling curve; required by the second-order path */
    st->xc = malloc(sizeof(double) * nc); st->yc = malloc(sizeof(double) * nc);
    RDYB_CHECK(st->xc && st->yc, RDYB_ERR_MEM, "out of memory");
    for (PetscInt c = 0; c < mesh->num_cells; ++c) { st->xc[c] = mesh->cells.centroids[c].X[0]; st->yc[c] = mesh->cells.centroids[c].X[1]; }
    v->cell_xc = st->xc; v->cell_yc = st->yc;
  }
  if (hr) {
    st->zc = malloc(sizeof(double) * nc);
    RDYB_CHECK(st->zc, RDYB_ERR_MEM, "out of memory");
    RDyCells    *cells    = &mesh->cells;
    RDyVertices *vertices = &mesh->vertices;
    for (PetscInt c = 0; c < mesh->num_cells; ++c) {
      if (config->grid.cell_elevation.file[0]) { /* cell elevation provided via a file */
        st->zc[c] = cells->centroids[c].X[2];
      } else { /* vertex-averaged bed elevation, summed in vertex order and divided once, as the reference does */
        PetscReal z_sum = 0.0;
        for (PetscInt k = cells->vertex_offsets[c]; k < cells->vertex_offsets[c + 1]; ++k) z_sum += vertices->points[cells->vertex_ids[k]].X[2];
        st->zc[c] = z_sum / (PetscReal)cells->num_vertices[c];
      }
    }
    v->cell_zc = st->zc;
  }
  if (second_order) {
    RDYB_CHECK(mesh->cells.centroids && mesh->edges.vertex_ids && mesh->vertices.points, RDYB_ERR_USER,
               "second-order reconstruction needs cell centroids and edge vertices in the mesh");
    st->xm = malloc(sizeof(double) * ne); st->ym = malloc(sizeof(double) * ne);
    RDYB_CHECK(st->xm && st->ym, RDYB_ERR_MEM, "out of memory");
    for (PetscInt e = 0; e < mesh->num_edges; ++e) {
      const PetscInt v0 = mesh->edges.vertex_ids[2 * e], v1 = mesh->edges.vertex_ids[2 * e + 1];
      st->xm[e] = 0.5 * (mesh->vertices.points[v0].X[0] + mesh->vertices.points[v1].X[0]);
      st->ym[e] = 0.5 * (mesh->vertices.points[v0].X[1] + mesh->vertices.points[v1].X[1]);
    }
    v->edge_xm = st->xm; v->edge_ym = st->ym;
  }
  return RDYB_SUCCESS;
}

static void FillConfig(const RDyConfig *config, int hr, RdybConfig *out) {
  memset(out, 0, sizeof(*out));
  out->tiny_h           = config->physics.flow.tiny_h;
  out->h_anuga_regular  = config->physics.flow.h_anuga_regular;
  out->xq2018_threshold = config->physics.flow.source.xq2018_threshold;
  out->source_method    = (int32_t)config->physics.flow.source.method;
  /* the factory that is called decides, as in CreateOperatorSubOperators: BS2002 takes the standard operators in the PETSc
   * backend (src/operator.c:202-209) */
  out->well_balancing = hr ? RDYB_WELL_BALANCING_HR : RDYB_WELL_BALANCING_NONE;
  out->temporal       = (int32_t)config->numerics.temporal;
  out->reconstruction = RDYB_RECON_FIRST_ORDER;
  if (config->numerics.second_order && !hr) { /* src/swe/swe_petsc.c:357-369 */
    const RDyLimiterType lim = config->numerics.no_limiter ? LIMITER_NONE : config->numerics.limiter;
    out->reconstruction = lim == LIMITER_NONE ? RDYB_RECON_MUSCL_NONE : lim == LIMITER_VANLEER ? RDYB_RECON_MUSCL_VANLEER : RDYB_RECON_MUSCL_MINMOD;
  }
}

static int adapter_device(void) {
  const char *env = getenv("RDYB_DEVICE"); /* one process per GPU: the launcher exports the local rank's device */
  return env ? atoi(env) : 0;
}

static PetscErrorCode create_flux(RDyConfig *config, RDyMesh *mesh, int hr, PetscInt num_boundaries, RDyBoundary *boundaries,
                                  RDyCondition *boundary_conditions, Vec *boundary_values, Vec *boundary_fluxes, Vec *boundary_fluxes_accum,
                                  OperatorDiagnostics *diagnostics, PetscOperator *flux_op) {
  RDYB_CHECK(config && mesh && flux_op, RDYB_ERR_ARG, "CreateB200FluxOperator: NULL argument");
  RDYB_CHECK(config->physics.flow.mode == FLOW_SWE, RDYB_ERR_SUP, "the B200 path implements the SWE flow mode only");
  RdybConfig cfg;
  FillConfig(config, hr, &cfg);
  MeshViewStorage st;
  int ierr = FillMeshView(config, mesh, hr, cfg.reconstruction != RDYB_RECON_FIRST_ORDER, &st);
  RdybBoundaryView *bv = calloc((size_t)(num_boundaries > 0 ? num_boundaries : 1), sizeof(*bv));
  if (!ierr && !bv) ierr = rdyb_fail(RDYB_ERR_MEM, "out of memory");
  for (PetscInt b = 0; !ierr && b < num_boundaries; ++b) {
    /* src/swe/swe_petsc.c:567-569 */
    if (!boundary_conditions[b].flow) { ierr = rdyb_fail(RDYB_ERR_USER, "Invalid boundary condition encountered for boundary %d\n", (int)boundaries[b].id); break; }
    bv[b].id        = (int32_t)boundaries[b].id;
    bv[b].index     = (int32_t)boundaries[b].index;
    bv[b].num_edges = (int32_t)boundaries[b].num_edges;
    bv[b].edge_ids  = (const int32_t *)boundaries[b].edge_ids;
    bv[b].bc_type   = (int32_t)boundary_conditions[b].flow->type;
  }
  B200Shared *shared = NULL;
  if (!ierr)
    ierr = b200_flux_create_core(&cfg, &st.view, (int32_t)num_boundaries, bv, boundary_values, boundary_fluxes, boundary_fluxes_accum, diagnostics,
                                 adapter_device(), &shared, flux_op);
  if (!ierr) ierr = adapter_register(mesh, shared);
  free(bv);
  FreeMeshView(&st); /* rdyb_operator_create copied what it needs to the device */
  return (PetscErrorCode)ierr;
}

static PetscErrorCode create_source(RDyConfig *config, RDyMesh *mesh, Vec external_sources, Vec material_properties, PetscOperator *source_op) {
  RDYB_CHECK(config && mesh && source_op, RDYB_ERR_ARG, "CreateB200SourceOperator: NULL argument");
  /* src/swe/swe_petsc.c:973-975 */
  RDYB_CHECK(config->physics.flow.source.method == SOURCE_SEMI_IMPLICIT || config->physics.flow.source.method == SOURCE_IMPLICIT_XQ2018, RDYB_ERR_USER,
             "Only semi_implicit and implicit_xq2018 are supported in the PETSc version");
  B200Shared *s = adapter_lookup(mesh);
  RDYB_CHECK(s, RDYB_ERR_USER, "CreateB200SourceOperator: create the flux operator of this mesh first (CreateOperatorSubOperators order, src/operator.c:195-216)");
  return (PetscErrorCode)b200_source_create_core(s, (int32_t)mesh->num_owned_cells, external_sources, material_properties, source_op);
}

PetscErrorCode CreateB200FluxOperator(RDyConfig *config, RDyMesh *mesh, MPI_Comm comm, PetscInt num_boundaries, RDyBoundary *boundaries,
                                      RDyCondition *boundary_conditions, Vec *boundary_values, Vec *boundary_fluxes, Vec *boundary_fluxes_accum,
                                      OperatorDiagnostics *diagnostics, PetscOperator *flux_op) {
  (void)comm; /* the halo of a partitioned mesh is set up on the device operator (RDyB200GetDeviceOperator) */
  return create_flux(config, mesh, 0, num_boundaries, boundaries, boundary_conditions, boundary_values, boundary_fluxes, boundary_fluxes_accum, diagnostics,
                     flux_op);
}
PetscErrorCode CreateB200FluxHROperator(RDyConfig *config, RDyMesh *mesh, PetscInt num_boundaries, RDyBoundary *boundaries,
                                        RDyCondition *boundary_conditions, Vec *boundary_values, Vec *boundary_fluxes, Vec *boundary_fluxes_accum,
                                        OperatorDiagnostics *diagnostics, PetscOperator *flux_op) {
  return create_flux(config, mesh, 1, num_boundaries, boundaries, boundary_conditions, boundary_values, boundary_fluxes, boundary_fluxes_accum, diagnostics,
                     flux_op);
}
PetscErrorCode CreateB200SourceOperator(RDyConfig *config, RDyMesh *mesh, Vec external_sources, Vec material_properties, PetscOperator *source_op) {
  return create_source(config, mesh, external_sources, material_properties, source_op);
}
PetscErrorCode CreateB200SourceHROperator(RDyConfig *config, RDyMesh *mesh, Vec external_sources, Vec material_properties, PetscOperator *source_op) {
  return create_source(config, mesh, external_sources, material_properties, source_op); /* the device operator already knows it is HR: no bed slope */
}

/* the device operator behind the sub-operators of `mesh` (kernel selection, halo setup, device-resident stepping); borrowed */
PetscErrorCode RDyB200GetDeviceOperator(RDyMesh *mesh, RdybOperator *device_op) {
  B200Shared *s = adapter_lookup(mesh);
  RDYB_CHECK(s && device_op, RDYB_ERR_USER, "no B200 operator was created on this mesh");
  *device_op = s->dev;
  return PETSC_SUCCESS;
}
/* fused = 1: flux.apply writes the complete right-hand side and source.apply does nothing (include/rdyb200_petsc.h) */
PetscErrorCode RDyB200SetFused(RDyMesh *mesh, int fused) {
  B200Shared *s = adapter_lookup(mesh);
  RDYB_CHECK(s, RDYB_ERR_USER, "no B200 operator was created on this mesh");
  s->fused = fused != 0; s->rhs_pending = 0;
  return PETSC_SUCCESS;
}

#ifdef RDYB_ADAPTER_REPLACE_PETSC
/* build RDycore WITHOUT src/operator_fluxes_petsc.c and src/operator_sources_petsc.c: the reference's own names */
PetscErrorCode CreatePetscFluxOperator(RDyConfig *c, RDyMesh *m, MPI_Comm comm, PetscInt nb, RDyBoundary *b, RDyCondition *bc, Vec *bv, Vec *bf, Vec *ba,
                                       OperatorDiagnostics *d, PetscOperator *op) { return CreateB200FluxOperator(c, m, comm, nb, b, bc, bv, bf, ba, d, op); }
PetscErrorCode CreatePetscFluxHROperator(RDyConfig *c, RDyMesh *m, PetscInt nb, RDyBoundary *b, RDyCondition *bc, Vec *bv, Vec *bf, Vec *ba,
                                         OperatorDiagnostics *d, PetscOperator *op) { return CreateB200FluxHROperator(c, m, nb, b, bc, bv, bf, ba, d, op); }
PetscErrorCode CreatePetscSourceOperator(RDyConfig *c, RDyMesh *m, Vec e, Vec p, PetscOperator *op) { return CreateB200SourceOperator(c, m, e, p, op); }
PetscErrorCode CreatePetscSourceHROperator(RDyConfig *c, RDyMesh *m, Vec e, Vec p, PetscOperator *op) { return CreateB200SourceHROperator(c, m, e, p, op); }
#endif
