/* petsc_operator_b200.c — the B200 SWE path behind RDycore's PetscOperator sub-operator interface (include/rdyb200_petsc.h).
 *
 * Host C only: every device action goes through the C-ABI of include/rdyb200.h.  Three parts:
 *   1. a host Vec stand-in for builds without PETSc;
 *   2. the PetscOperator object (single / composite, named fields) with the semantics of src/petsc_operator.c;
 *   3. the flux and source factories whose apply callbacks drive ONE fused kernel run per stage.
 */
#include "rdyb200_petsc.h"

#include <stdlib.h>
#include <string.h>

#include "rdyb_error.h"

/* ------------------------------------------------------------------ 1. host Vec ---- */
#ifndef RDYB_HAVE_PETSC
struct RdybVec_ {
  double *array;
  int64_t n;
  int32_t bs;
  int     pinned;
  int64_t state; /* bumped by every restore_array / zero / copy-into, like PetscObjectStateIncrease */
  int     out;   /* arrays currently handed out */
};

int rdyb_vec_create(int64_t n, int32_t bs, int pinned, Vec *v) {
  RDYB_CHECK(v, RDYB_ERR_ARG, "rdyb_vec_create: NULL output");
  RDYB_CHECK(n >= 0 && bs >= 1 && n % bs == 0, RDYB_ERR_ARG_SIZ, "Local size %lld not compatible with block size %d", (long long)n, bs);
  struct RdybVec_ *x = calloc(1, sizeof(*x));
  RDYB_CHECK(x, RDYB_ERR_MEM, "out of memory");
  const size_t bytes = sizeof(double) * (size_t)(n > 0 ? n : 1);
  if (pinned) {
    void *p   = NULL;
    int  ierr = rdyb_host_alloc(bytes, &p); /* fails loudly without a GPU */
    if (ierr) { free(x); return ierr; }
    x->array = p;
    memset(x->array, 0, bytes);
  } else {
    x->array = calloc(1, bytes);
    if (!x->array) { free(x); return rdyb_fail(RDYB_ERR_MEM, "out of memory (%zu bytes)", bytes); }
  }
  x->n = n; x->bs = bs; x->pinned = pinned; x->state = 1;
  *v = x;
  return RDYB_SUCCESS;
}

int rdyb_vec_destroy(Vec *v) {
  if (!v || !*v) return RDYB_SUCCESS;
  if ((*v)->pinned) rdyb_host_free((*v)->array);
  else free((*v)->array);
  free(*v);
  *v = NULL;
  return RDYB_SUCCESS;
}

int rdyb_vec_get_array(Vec v, double **array) {
  RDYB_CHECK(v && array, RDYB_ERR_ARG, "VecGetArray: NULL argument");
  ++v->out;
  *array = v->array;
  return RDYB_SUCCESS;
}

int rdyb_vec_restore_array(Vec v, double **array) {
  RDYB_CHECK(v, RDYB_ERR_ARG, "VecRestoreArray: NULL Vec");
  RDYB_CHECK(v->out > 0, RDYB_ERR_USER, "VecRestoreArray without a matching VecGetArray");
  --v->out;
  ++v->state;
  if (array) *array = NULL;
  return RDYB_SUCCESS;
}

int rdyb_vec_get_local_size(Vec v, int64_t *n) {
  RDYB_CHECK(v && n, RDYB_ERR_ARG, "VecGetLocalSize: NULL argument");
  *n = v->n;
  return RDYB_SUCCESS;
}

int rdyb_vec_get_block_size(Vec v, int32_t *bs) {
  RDYB_CHECK(v && bs, RDYB_ERR_ARG, "VecGetBlockSize: NULL argument");
  *bs = v->bs;
  return RDYB_SUCCESS;
}

int rdyb_vec_get_state(Vec v, int64_t *state) {
  RDYB_CHECK(v && state, RDYB_ERR_ARG, "PetscObjectStateGet: NULL argument");
  *state = v->state;
  return RDYB_SUCCESS;
}

int rdyb_vec_zero(Vec v) {
  RDYB_CHECK(v, RDYB_ERR_ARG, "VecZeroEntries: NULL Vec");
  memset(v->array, 0, sizeof(double) * (size_t)v->n);
  ++v->state;
  return RDYB_SUCCESS;
}

int rdyb_vec_copy(Vec x, Vec y) {
  RDYB_CHECK(x && y, RDYB_ERR_ARG, "VecCopy: NULL Vec");
  RDYB_CHECK(x->n == y->n, RDYB_ERR_ARG_SIZ, "Incompatible vector local lengths parameter # 1 local size %lld != parameter # 2 local size %lld",
             (long long)x->n, (long long)y->n);
  if (x != y) memcpy(y->array, x->array, sizeof(double) * (size_t)x->n);
  ++y->state;
  return RDYB_SUCCESS;
}

/* the few Vec calls the sub-operators below make, spelled as in PETSc */
static int VecGetArray(Vec v, double **a) { return rdyb_vec_get_array(v, a); }
static int VecRestoreArray(Vec v, double **a) { return rdyb_vec_restore_array(v, a); }
static int VecGetBlockSize(Vec v, PetscInt *bs) { return rdyb_vec_get_block_size(v, bs); }
static int VecLocalSize64(Vec v, int64_t *n) { return rdyb_vec_get_local_size(v, n); }
static int VecState64(Vec v, int64_t *s) { return rdyb_vec_get_state(v, s); }
#else
static int VecLocalSize64(Vec v, int64_t *n) { PetscInt m; PetscErrorCode e = VecGetLocalSize(v, &m); *n = m; return (int)e; }
static int VecState64(Vec v, int64_t *s) { PetscObjectState t; PetscErrorCode e = PetscObjectStateGet((PetscObject)v, &t); *s = (int64_t)t; return (int)e; }
#endif

/* ------------------------------------------------------------------ 2. PetscOperator ---- */
typedef struct {
  PetscInt           count, capacity;
  RdybPetscOperator *subs;
} CompositeContext;

int rdyb_petscop_create(void *context, PetscErrorCode (*apply)(void *, RdybPetscOperatorFields, PetscReal, Vec, Vec),
                        PetscErrorCode (*destroy)(void *), RdybPetscOperator *op) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "PetscOperatorCreate: NULL output");
  RDYB_CHECK(apply && destroy, RDYB_ERR_ARG, "PetscOperatorCreate: apply and destroy are required");
  RdybPetscOperator o = calloc(1, sizeof(*o));
  RDYB_CHECK(o, RDYB_ERR_MEM, "out of memory");
  o->context = context; o->apply = apply; o->destroy = destroy; o->is_composite = PETSC_FALSE;
  *op = o;
  return RDYB_SUCCESS;
}

int rdyb_petscop_destroy(RdybPetscOperator *op) {
  if (!op || !*op) return RDYB_SUCCESS;
  RdybPetscOperator o = *op;
  free(o->fields.fields); /* the Vecs behind the fields are not ours */
  int ierr = o->destroy(o->context);
  free(o);
  *op = NULL;
  return ierr;
}

int rdyb_petscop_apply(RdybPetscOperator op, PetscReal dt, Vec u_local, Vec f_global) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "PetscOperatorApply: NULL operator");
  return op->apply(op->context, op->fields, dt, u_local, f_global);
}

static int composite_apply(void *context, RdybPetscOperatorFields fields, PetscReal dt, Vec u_local, Vec f_global) {
  (void)fields; /* every sub-operator carries its own copy of the fields (set_field propagates) */
  CompositeContext *c = context;
  for (PetscInt i = 0; i < c->count; ++i) RDYB_CALL(rdyb_petscop_apply(c->subs[i], dt, u_local, f_global)); /* insertion order */
  return RDYB_SUCCESS;
}

static int composite_destroy(void *context) {
  CompositeContext *c = context;
  int first = RDYB_SUCCESS;
  for (PetscInt i = 0; i < c->count; ++i) {
    int ierr = rdyb_petscop_destroy(&c->subs[i]);
    if (ierr && !first) first = ierr;
  }
  free(c->subs);
  free(c);
  return first;
}

int rdyb_petscop_create_composite(RdybPetscOperator *op) {
  CompositeContext *c = calloc(1, sizeof(*c));
  RDYB_CHECK(c, RDYB_ERR_MEM, "out of memory");
  int ierr = rdyb_petscop_create(c, composite_apply, composite_destroy, op);
  if (ierr) { free(c); return ierr; }
  (*op)->is_composite = PETSC_TRUE;
  return RDYB_SUCCESS;
}

int rdyb_petscop_composite_add_sub(RdybPetscOperator op, RdybPetscOperator sub_op) {
  RDYB_CHECK(op && sub_op, RDYB_ERR_ARG, "PetscOperatorCompositeAddSub: NULL operator");
  RDYB_CHECK(op->is_composite, RDYB_ERR_ARG, "PetscOperatorCompositeAddSub: not a composite operator");
  CompositeContext *c = op->context;
  if (c->count == c->capacity) {
    const PetscInt cap = c->capacity ? 2 * c->capacity : 8;
    RdybPetscOperator *s = realloc(c->subs, sizeof(*s) * (size_t)cap);
    RDYB_CHECK(s, RDYB_ERR_MEM, "out of memory");
    c->subs = s; c->capacity = cap;
  }
  c->subs[c->count++] = sub_op;
  return RDYB_SUCCESS;
}

int rdyb_petscop_fields_get(RdybPetscOperatorFields fields, const char *name, Vec *vec) {
  RDYB_CHECK(name && vec, RDYB_ERR_ARG, "PetscOperatorFieldsGet: NULL argument");
  *vec = NULL;
  for (PetscInt f = 0; f < fields.num_fields; ++f)
    if (!strncmp(name, fields.fields[f].name, RDYB_MAX_NAME_LEN)) { *vec = fields.fields[f].vec; break; }
  return RDYB_SUCCESS;
}

int rdyb_petscop_set_field(RdybPetscOperator op, const char *name, Vec vec) {
  RDYB_CHECK(op && name, RDYB_ERR_ARG, "PetscOperatorSetField: NULL argument");
  RdybPetscOperatorFields *F = &op->fields;
  PetscInt at = -1;
  for (PetscInt f = 0; f < F->num_fields; ++f)
    if (!strncmp(name, F->fields[f].name, RDYB_MAX_NAME_LEN)) at = f;
  if (at < 0) { /* a new name: append */
    if (F->num_fields == F->capacity) {
      const PetscInt cap = F->capacity ? 2 * F->capacity : 8;
      RdybPetscOperatorField *p = realloc(F->fields, sizeof(*p) * (size_t)cap);
      RDYB_CHECK(p, RDYB_ERR_MEM, "out of memory");
      F->fields = p; F->capacity = cap;
    }
    at = F->num_fields++;
    memset(&F->fields[at], 0, sizeof(F->fields[at]));
    strncpy(F->fields[at].name, name, RDYB_MAX_NAME_LEN);
  }
  F->fields[at].vec = vec;
  if (op->is_composite) {
    CompositeContext *c = op->context;
    for (PetscInt i = 0; i < c->count; ++i) RDYB_CALL(rdyb_petscop_set_field(c->subs[i], name, vec));
  }
  return RDYB_SUCCESS;
}

/* ------------------------------------------------------------------ 3. the B200 flux and source sub-operators ---- */
/* One GPU operator shared by the flux and the source sub-operator of an Operator (reference-counted: RDycore destroys
 * its flux and source composites independently, src/operator.c:459-472). */
typedef struct {
  RdybOperator dev;
  int          refs;
  int32_t      n_owned, n_view_cells;
  /* boundaries (borrowed Vec arrays, src/operator.c:368-379) */
  int32_t  nb;
  int32_t *b_edges, *b_type;
  Vec     *b_values, *b_fluxes, *b_accum;
  int64_t *b_values_state;
  RdybCourantDiagnostics *diags;
  /* source side (borrowed Vecs, src/operator.c:93-96) */
  int     has_source;
  Vec     ext_src, mat_props;
  int64_t ext_state, mat_state;
  /* the complete right-hand side of the last flux.apply, waiting for source.apply */
  double *rhs;      /* [3*n_owned], page-locked */
  double *bscratch; /* [3*max boundary edges] */
  int     rhs_pending, fused, want_prim;
} B200Shared;

static int shared_release(B200Shared *s) {
  if (--s->refs > 0) return RDYB_SUCCESS;
  int ierr = rdyb_operator_destroy(&s->dev);
  if (s->rhs) rdyb_host_free(s->rhs);
  free(s->bscratch); free(s->b_edges); free(s->b_type); free(s->b_values_state);
  free(s);
  return ierr;
}

/* push host-side operator data that changed since the last apply to the device */
static int shared_refresh_inputs(B200Shared *s) {
  for (int32_t b = 0; b < s->nb; ++b) {
    if (s->b_type[b] != RDYB_CONDITION_DIRICHLET || !s->b_values || !s->b_values[b] || s->b_edges[b] == 0) continue;
    int64_t st;
    RDYB_CALL(VecState64(s->b_values[b], &st));
    if (st == s->b_values_state[b]) continue;
    double *v;
    RDYB_CALL(VecGetArray(s->b_values[b], &v));
    int ierr = rdyb_operator_set_boundary_values(s->dev, b, 0, 3, s->b_edges[b], v); /* values[3*e + c], as the Vec stores them */
    RDYB_CALL(VecRestoreArray(s->b_values[b], &v));
    RDYB_CALL(ierr);
    RDYB_CALL(VecState64(s->b_values[b], &s->b_values_state[b])); /* our own restore bumped it */
  }
  if (s->has_source) {
    int64_t st;
    RDYB_CALL(VecState64(s->ext_src, &st));
    if (st != s->ext_state) {
      double *v;
      RDYB_CALL(VecGetArray(s->ext_src, &v));
      int ierr = rdyb_operator_set_external_source_aos(s->dev, s->n_owned, v);
      RDYB_CALL(VecRestoreArray(s->ext_src, &v));
      RDYB_CALL(ierr);
      RDYB_CALL(VecState64(s->ext_src, &s->ext_state));
    }
    RDYB_CALL(VecState64(s->mat_props, &st));
    if (st != s->mat_state) {
      double *v;
      RDYB_CALL(VecGetArray(s->mat_props, &v));
      int ierr = rdyb_operator_set_mannings(s->dev, s->n_owned, v);
      RDYB_CALL(VecRestoreArray(s->mat_props, &v));
      RDYB_CALL(ierr);
      RDYB_CALL(VecState64(s->mat_props, &s->mat_state));
    }
  }
  return RDYB_SUCCESS;
}

/* boundary_fluxes <- this apply's edge fluxes; boundary_fluxes_accum += dt * boundary_fluxes (src/swe/swe_petsc.c:574-623) */
static int shared_collect_boundary_fluxes(B200Shared *s, double dt) {
  if (!s->b_fluxes) return RDYB_SUCCESS;
  for (int32_t b = 0; b < s->nb; ++b) {
    const int32_t ne = s->b_edges[b];
    if (ne == 0 || !s->b_fluxes[b]) continue;
    RDYB_CALL(rdyb_operator_get_boundary_fluxes(s->dev, b, ne, 0, s->bscratch)); /* [c][e] */
    double *f, *acc = NULL;
    RDYB_CALL(VecGetArray(s->b_fluxes[b], &f));
    if (s->b_accum && s->b_accum[b]) RDYB_CALL(VecGetArray(s->b_accum[b], &acc));
    for (int32_t e = 0; e < ne; ++e)
      for (int c = 0; c < 3; ++c) {
        const double v = s->bscratch[(size_t)c * ne + e];
        f[3 * e + c] = v;
        if (acc) acc[3 * e + c] += dt * v;
      }
    RDYB_CALL(VecRestoreArray(s->b_fluxes[b], &f));
    if (acc) RDYB_CALL(VecRestoreArray(s->b_accum[b], &acc));
  }
  return RDYB_SUCCESS;
}

static int check_apply_vectors(const B200Shared *s, Vec u_local, Vec f_global) {
  RDYB_CHECK(u_local && f_global, RDYB_ERR_ARG, "apply: NULL Vec");
  PetscInt n_dof;
  RDYB_CALL(VecGetBlockSize(u_local, &n_dof));
  RDYB_CHECK(n_dof == 3, RDYB_ERR_USER, "Number of dof in local vector must be 3!"); /* src/swe/swe_petsc.c:231-233 */
  int64_t nu, nf;
  RDYB_CALL(VecLocalSize64(u_local, &nu));
  RDYB_CALL(VecLocalSize64(f_global, &nf));
  RDYB_CHECK(nu == 3 * (int64_t)s->n_view_cells, RDYB_ERR_ARG_SIZ, "local vector has %lld entries, the mesh has %d cells x 3", (long long)nu, s->n_view_cells);
  RDYB_CHECK(nf == 3 * (int64_t)s->n_owned, RDYB_ERR_ARG_SIZ, "global vector has %lld entries, the mesh has %d owned cells x 3", (long long)nf, s->n_owned);
  return RDYB_SUCCESS;
}

/* ApplyInteriorFlux + ApplyBoundaryFlux x B (+ the sources, fused): PetscOperator apply shape */
static PetscErrorCode b200_flux_apply(void *context, RdybPetscOperatorFields fields, PetscReal dt, Vec u_local, Vec f_global) {
  (void)fields;
  B200Shared *s = context;
  RDYB_CALL(check_apply_vectors(s, u_local, f_global));
  RDYB_CALL(shared_refresh_inputs(s));
  /* courant on (the reference always computes it), flux divergence only when the split is visible, boundary fluxes when the Vecs exist */
  RDYB_CALL(rdyb_operator_set_options(s->dev, 1, !s->fused, s->want_prim, s->b_fluxes != NULL && s->nb > 0));
  RDYB_CALL(rdyb_operator_reset_diagnostics(s->dev));
  double *u, *f;
  RDYB_CALL(VecGetArray(u_local, &u));
  RDYB_CALL(VecGetArray(f_global, &f));
  int ierr;
  if (s->fused) {
    ierr = rdyb_operator_apply_host(s->dev, dt, u, f); /* complete right-hand side straight into F */
  } else {
    ierr = rdyb_operator_apply_host(s->dev, dt, u, s->rhs);
    if (!ierr) ierr = rdyb_operator_get_field(s->dev, 0, f); /* F <- flux divergence (F was zero: VecZeroEntries, src/rdysetup.c:1130) */
    s->rhs_pending = !ierr;
  }
  RDYB_CALL(VecRestoreArray(u_local, &u));
  RDYB_CALL(VecRestoreArray(f_global, &f));
  RDYB_CALL(ierr);
  RDYB_CALL(shared_collect_boundary_fluxes(s, dt));
  if (s->diags) { /* running max, first-seen wins ties (src/swe/swe_petsc.c:291) */
    RdybCourantDiagnostics d;
    RDYB_CALL(rdyb_operator_get_courant(s->dev, &d));
    if (d.max_courant_num > s->diags->max_courant_num) *s->diags = d;
  }
  return PETSC_SUCCESS;
}

static PetscErrorCode b200_flux_destroy(void *context) { return shared_release(context); }

/* ApplySourceSemiImplicit / ApplySourceImplicitXQ2018: PetscOperator apply shape.  The arithmetic already ran inside the
 * fused kernel of flux.apply; this hands its result over. */
static PetscErrorCode b200_source_apply(void *context, RdybPetscOperatorFields fields, PetscReal dt, Vec u_local, Vec f_global) {
  B200Shared *s = context;
  RDYB_CALL(check_apply_vectors(s, u_local, f_global));
  Vec prim = NULL;
  RDYB_CALL(rdyb_petscop_fields_get(fields, "primitive_variables", &prim));
  if (!s->fused) {
    RDYB_CHECK(s->rhs_pending, RDYB_ERR_USER,
               "the B200 source operator must be applied right after its flux operator (ApplyPetscOperator order, src/operator.c:660-666)");
    double *f;
    RDYB_CALL(VecGetArray(f_global, &f));
    memcpy(f, s->rhs, sizeof(double) * 3 * (size_t)s->n_owned);
    RDYB_CALL(VecRestoreArray(f_global, &f));
    s->rhs_pending = 0;
  }
  if (prim) {
    if (!s->want_prim) {
      /* first apply after the field was set: flux.apply did not know yet.  The device still holds this u_local, so run the
       * kernel once more with the output switched on; from now on flux.apply produces it in its one run. */
      s->want_prim = 1;
      RDYB_CALL(rdyb_operator_set_options(s->dev, 0, 0, 1, 0)); /* nothing else is recorded twice */
      RDYB_CALL(rdyb_operator_apply(s->dev, dt, NULL));
    }
    double *p;
    RDYB_CALL(VecGetArray(prim, &p));
    int ierr = rdyb_operator_get_field(s->dev, 1, p);
    RDYB_CALL(VecRestoreArray(prim, &p));
    RDYB_CALL(ierr);
  }
  return PETSC_SUCCESS;
}

static PetscErrorCode b200_source_destroy(void *context) { return shared_release(context); }

int rdyb_create_flux_operator(const RdybConfig *config, const RdybMeshView *mesh, int32_t nb, const RdybBoundaryView *boundaries, Vec *boundary_values,
                              Vec *boundary_fluxes, Vec *boundary_fluxes_accum, RdybCourantDiagnostics *diagnostics, int device,
                              RdybPetscOperator *flux_op) {
  RDYB_CHECK(config && mesh && flux_op, RDYB_ERR_ARG, "rdyb_create_flux_operator: NULL argument");
  B200Shared *s = calloc(1, sizeof(*s));
  RDYB_CHECK(s, RDYB_ERR_MEM, "out of memory");
  s->refs = 1;
  int ierr = rdyb_operator_create(config, mesh, nb, boundaries, NULL, device, NULL, &s->dev);
  if (ierr) { free(s); return ierr; }
  s->n_owned = mesh->num_owned_cells; s->n_view_cells = mesh->num_cells;
  s->nb = nb; s->b_values = boundary_values; s->b_fluxes = boundary_fluxes; s->b_accum = boundary_fluxes_accum; s->diags = diagnostics;
  s->b_edges = calloc((size_t)nb + 1, sizeof(int32_t)); s->b_type = calloc((size_t)nb + 1, sizeof(int32_t));
  s->b_values_state = calloc((size_t)nb + 1, sizeof(int64_t));
  int32_t maxe = 1;
  for (int32_t b = 0; b < nb; ++b) {
    s->b_edges[b] = boundaries[b].num_edges; s->b_type[b] = boundaries[b].bc_type;
    if (boundaries[b].num_edges > maxe) maxe = boundaries[b].num_edges;
  }
  s->bscratch = malloc(sizeof(double) * 3 * (size_t)maxe);
  void *p = NULL;
  ierr = rdyb_host_alloc(sizeof(double) * 3 * (size_t)(s->n_owned > 0 ? s->n_owned : 1), &p);
  s->rhs = p;
  RdybPetscOperator sub = NULL;
  if (!ierr) ierr = rdyb_petscop_create_composite(flux_op); /* a composite, as CreatePetscFluxOperator returns (src/operator_fluxes_petsc.c:24) */
  if (!ierr) ierr = rdyb_petscop_create(s, b200_flux_apply, b200_flux_destroy, &sub);
  if (!ierr) ierr = rdyb_petscop_composite_add_sub(*flux_op, sub);
  if (ierr) { shared_release(s); return ierr; }
  return RDYB_SUCCESS;
}

static B200Shared *shared_of(RdybPetscOperator op) {
  if (!op) return NULL;
  if (op->is_composite) {
    CompositeContext *c = op->context;
    for (PetscInt i = 0; i < c->count; ++i) {
      B200Shared *s = shared_of(c->subs[i]);
      if (s) return s;
    }
    return NULL;
  }
  return (op->apply == b200_flux_apply || op->apply == b200_source_apply) ? op->context : NULL;
}

int rdyb_create_source_operator(const RdybConfig *config, const RdybMeshView *mesh, Vec external_sources, Vec material_properties,
                                RdybPetscOperator flux_op, RdybPetscOperator *source_op) {
  RDYB_CHECK(config && mesh && source_op, RDYB_ERR_ARG, "rdyb_create_source_operator: NULL argument");
  B200Shared *s = shared_of(flux_op);
  RDYB_CHECK(s, RDYB_ERR_USER, "rdyb_create_source_operator needs the flux operator made by rdyb_create_flux_operator (they share one GPU operator)");
  RDYB_CHECK(mesh->num_owned_cells == s->n_owned, RDYB_ERR_ARG_SIZ, "flux and source operators are built on different meshes");
  RDYB_CHECK(external_sources && material_properties, RDYB_ERR_ARG, "rdyb_create_source_operator: NULL Vec");
  int64_t n;
  RDYB_CALL(VecLocalSize64(external_sources, &n));
  RDYB_CHECK(n == 3 * (int64_t)s->n_owned, RDYB_ERR_ARG_SIZ, "external_sources has %lld entries (should be %d x 3)", (long long)n, s->n_owned);
  RDYB_CALL(VecLocalSize64(material_properties, &n));
  RDYB_CHECK(n == (int64_t)s->n_owned, RDYB_ERR_ARG_SIZ, "material_properties has %lld entries (should be %d)", (long long)n, s->n_owned);
  s->has_source = 1; s->ext_src = external_sources; s->mat_props = material_properties; s->ext_state = s->mat_state = -1;
  RdybPetscOperator sub = NULL;
  RDYB_CALL(rdyb_petscop_create_composite(source_op)); /* src/operator_sources_petsc.c:19 */
  int ierr = rdyb_petscop_create(s, b200_source_apply, b200_source_destroy, &sub);
  if (!ierr) ierr = rdyb_petscop_composite_add_sub(*source_op, sub);
  if (ierr) { rdyb_petscop_destroy(source_op); return ierr; }
  ++s->refs;
  return RDYB_SUCCESS;
}

int rdyb_petscop_set_fused(RdybPetscOperator flux_op, int fused) {
  B200Shared *s = shared_of(flux_op);
  RDYB_CHECK(s, RDYB_ERR_USER, "not a B200 flux operator");
  s->fused = fused != 0; s->rhs_pending = 0;
  return RDYB_SUCCESS;
}

int rdyb_petscop_get_device_operator(RdybPetscOperator op, RdybOperator *device_op) {
  B200Shared *s = shared_of(op);
  RDYB_CHECK(s && device_op, RDYB_ERR_USER, "not a B200 flux or source operator");
  *device_op = s->dev;
  return RDYB_SUCCESS;
}
