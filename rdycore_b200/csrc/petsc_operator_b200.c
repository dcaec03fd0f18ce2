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
/* the sub-operators themselves live in b200_suboperators.h, shared with the RDycore adapter (rdycore_adapter.c); here they are
 * bound to the stand-in PetscOperator / Vec of this file */
#define B200_POP RdybPetscOperator
#define B200_POPFIELDS RdybPetscOperatorFields
#define B200_POP_CREATE rdyb_petscop_create
#define B200_POP_DESTROY rdyb_petscop_destroy
#define B200_POP_CREATE_COMPOSITE rdyb_petscop_create_composite
#define B200_POP_ADD_SUB rdyb_petscop_composite_add_sub
#define B200_POP_FIELDS_GET rdyb_petscop_fields_get
#define B200_DIAGS_T RdybCourantDiagnostics
#define B200_DIAGS_STORE(ptr, d)                                                                                        \
  do { /* running max, first-seen wins ties (src/swe/swe_petsc.c:291) */                                                \
    if ((d).max_courant_num > (ptr)->max_courant_num) *(ptr) = (d);                                                     \
  } while (0)
#define B200_ON_RELEASE(s) ((void)(s))
#include "b200_suboperators.h"

int rdyb_create_flux_operator(const RdybConfig *config, const RdybMeshView *mesh, int32_t nb, const RdybBoundaryView *boundaries, Vec *boundary_values,
                              Vec *boundary_fluxes, Vec *boundary_fluxes_accum, RdybCourantDiagnostics *diagnostics, int device,
                              RdybPetscOperator *flux_op) {
  return b200_flux_create_core(config, mesh, nb, boundaries, boundary_values, boundary_fluxes, boundary_fluxes_accum, diagnostics, device, NULL, flux_op);
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
  return b200_source_create_core(s, mesh->num_owned_cells, external_sources, material_properties, source_op);
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
