/* b200_suboperators.h — the B200 flux and source sub-operators behind RDycore's PetscOperator interface, written once against
 * a handful of macros so that the SAME code is compiled
 *   (a) into librdyb200.so with the stand-in Vec / PetscOperator of include/rdyb200_petsc.h (petsc_operator_b200.c), and
 *   (b) into an RDycore build with the reference's own types - PetscOperator, PetscOperatorFields, Vec, OperatorDiagnostics of
 *       include/private/rdyoperatorimpl.h:21-90 - by rdycore_adapter.c, whose factories carry the reference's signatures
 *       (include/private/rdyoperatorimpl.h:234-239).
 * The includer defines, before including this file:
 *   B200_POP, B200_POPFIELDS                      operator handle and fields types
 *   B200_POP_CREATE(ctx, apply, destroy, &op)     PetscOperatorCreate
 *   B200_POP_DESTROY(&op)                         PetscOperatorDestroy
 *   B200_POP_CREATE_COMPOSITE(&op)                PetscOperatorCreateComposite
 *   B200_POP_ADD_SUB(op, sub)                     PetscOperatorCompositeAddSub
 *   B200_POP_FIELDS_GET(fields, name, &vec)       PetscOperatorFieldsGet
 *   VecGetArray / VecRestoreArray / VecGetBlockSize, VecLocalSize64(v, &n), VecState64(v, &state)
 *   B200_DIAGS_T, B200_DIAGS_STORE(ptr, d)        where the running-max Courant number goes (d is an RdybCourantDiagnostics)
 *   B200_ON_RELEASE(shared)                       hook run when the last sub-operator of a pair is destroyed
 * Split at the ABI, fused underneath (include/rdyb200_petsc.h explains the contract of the two apply callbacks).
 */
#ifndef RDYB_B200_SUBOPERATORS_H
#define RDYB_B200_SUBOPERATORS_H
#include <stdlib.h>
#include <string.h>

#include "rdyb200.h"
#include "rdyb_error.h"

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
  B200_DIAGS_T *diags;
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
  B200_ON_RELEASE(s);
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
static PetscErrorCode b200_flux_apply(void *context, B200_POPFIELDS fields, PetscReal dt, Vec u_local, Vec f_global) {
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
    B200_DIAGS_STORE(s->diags, d);
  }
  return PETSC_SUCCESS;
}

static PetscErrorCode b200_flux_destroy(void *context) { return shared_release(context); }

/* ApplySourceSemiImplicit / ApplySourceImplicitXQ2018: PetscOperator apply shape.  The arithmetic already ran inside the
 * fused kernel of flux.apply; this hands its result over. */
static PetscErrorCode b200_source_apply(void *context, B200_POPFIELDS fields, PetscReal dt, Vec u_local, Vec f_global) {
  B200Shared *s = context;
  RDYB_CALL(check_apply_vectors(s, u_local, f_global));
  Vec prim = NULL;
  RDYB_CALL(B200_POP_FIELDS_GET(fields, "primitive_variables", &prim));
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


/* CreatePetscFlux[HR]Operator: one GPU operator + a composite holding ONE sub-operator (src/operator_fluxes_petsc.c:17-95) */
static int b200_flux_create_core(const RdybConfig *config, const RdybMeshView *mesh, int32_t nb, const RdybBoundaryView *boundaries, Vec *boundary_values,
                                 Vec *boundary_fluxes, Vec *boundary_fluxes_accum, B200_DIAGS_T *diagnostics, int device, B200Shared **shared,
                                 B200_POP *flux_op) {
  RDYB_CHECK(config && mesh && flux_op, RDYB_ERR_ARG, "B200 flux operator: NULL argument");
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
    s->b_edges[b] = boundaries[b].num_edges; s->b_type[b] = boundaries[b].bc_type; s->b_values_state[b] = -1;
    if (boundaries[b].num_edges > maxe) maxe = boundaries[b].num_edges;
  }
  s->bscratch = malloc(sizeof(double) * 3 * (size_t)maxe);
  void *p = NULL;
  ierr = rdyb_host_alloc(sizeof(double) * 3 * (size_t)(s->n_owned > 0 ? s->n_owned : 1), &p);
  s->rhs = p;
  B200_POP sub = NULL;
  if (!ierr) ierr = B200_POP_CREATE_COMPOSITE(flux_op); /* a composite, as CreatePetscFluxOperator returns (src/operator_fluxes_petsc.c:24) */
  if (!ierr) ierr = B200_POP_CREATE(s, b200_flux_apply, b200_flux_destroy, &sub);
  if (!ierr) ierr = B200_POP_ADD_SUB(*flux_op, sub);
  if (ierr) { shared_release(s); return ierr; }
  if (shared) *shared = s;
  return RDYB_SUCCESS;
}

/* CreatePetscSource[HR]Operator (src/operator_sources_petsc.c:13-45): the second face of the same GPU operator */
static int b200_source_create_core(B200Shared *s, int32_t num_owned, Vec external_sources, Vec material_properties, B200_POP *source_op) {
  RDYB_CHECK(s && source_op, RDYB_ERR_ARG, "B200 source operator: NULL argument");
  RDYB_CHECK(num_owned == s->n_owned, RDYB_ERR_ARG_SIZ, "flux and source operators are built on different meshes");
  RDYB_CHECK(external_sources && material_properties, RDYB_ERR_ARG, "B200 source operator: NULL Vec");
  int64_t n;
  RDYB_CALL(VecLocalSize64(external_sources, &n));
  RDYB_CHECK(n == 3 * (int64_t)s->n_owned, RDYB_ERR_ARG_SIZ, "external_sources has %lld entries (should be %d x 3)", (long long)n, s->n_owned);
  RDYB_CALL(VecLocalSize64(material_properties, &n));
  RDYB_CHECK(n == (int64_t)s->n_owned, RDYB_ERR_ARG_SIZ, "material_properties has %lld entries (should be %d)", (long long)n, s->n_owned);
  s->has_source = 1; s->ext_src = external_sources; s->mat_props = material_properties; s->ext_state = s->mat_state = -1;
  B200_POP sub = NULL;
  RDYB_CALL(B200_POP_CREATE_COMPOSITE(source_op)); /* src/operator_sources_petsc.c:19 */
  int ierr = B200_POP_CREATE(s, b200_source_apply, b200_source_destroy, &sub);
  if (!ierr) ierr = B200_POP_ADD_SUB(*source_op, sub);
  if (ierr) { B200_POP_DESTROY(source_op); return ierr; }
  ++s->refs;
  return RDYB_SUCCESS;
}
#endif /* RDYB_B200_SUBOPERATORS_H */
