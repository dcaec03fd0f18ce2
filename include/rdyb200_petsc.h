/* rdyb200_petsc.h — the B200 SWE path behind RDycore's PetscOperator sub-operator interface.
 *
 * RDycore builds its right-hand side from two composite PetscOperators, "flux" and "source"
 * (CreateOperatorSubOperators, src/operator.c:195-216), applies them back to back
 * (ApplyPetscOperator, src/operator.c:656-672) and hands named Vecs ("riemannf", "primitive_variables")
 * to the source operator (src/operator.c:224-299).  This header declares
 *   (1) the PetscOperator object itself, with the layout and semantics of
 *       include/private/rdyoperatorimpl.h:45-90 and src/petsc_operator.c:11-166, for builds where RDycore is absent;
 *   (2) the two factories that replace CreatePetscFlux[HR]Operator / CreatePetscSource[HR]Operator
 *       (include/private/rdyoperatorimpl.h:234-239; src/operator_fluxes_petsc.c:17-95, src/operator_sources_petsc.c:13-45):
 *       their apply/destroy callbacks have exactly the PetscOperator shape and their context is the GPU operator handle.
 *
 * Vec: with -DRDYB_HAVE_PETSC the real <petscvec.h> types are used.  Otherwise `Vec` is the small host vector below
 * (array + local size + block size), enough for VecGetArray/VecRestoreArray/VecGetBlockSize/VecGetLocalSize — the only
 * Vec calls the reference's sub-operators make (src/swe/swe_petsc.c:228-243, 312-315, 520-530).
 */
#ifndef RDYB200_PETSC_H
#define RDYB200_PETSC_H

#include "rdyb200.h"

#ifdef __cplusplus
extern "C" {
#endif

#ifdef RDYB_HAVE_PETSC
#include <petscvec.h>
#else
typedef int     PetscErrorCode;
typedef double  PetscReal;
typedef int32_t PetscInt;
typedef int     PetscBool;
#define PETSC_SUCCESS 0
#define PETSC_TRUE 1
#define PETSC_FALSE 0
typedef struct RdybVec_ *Vec;
/* host Vec stand-in; `pinned` != 0 allocates page-locked memory (faster host<->device copies) */
int rdyb_vec_create(int64_t local_size, int32_t block_size, int pinned, Vec *v);
int rdyb_vec_destroy(Vec *v);
int rdyb_vec_get_array(Vec v, double **array);      /* VecGetArray */
int rdyb_vec_restore_array(Vec v, double **array);  /* VecRestoreArray: bumps the Vec's state counter like PetscObjectStateIncrease */
int rdyb_vec_get_local_size(Vec v, int64_t *n);     /* VecGetLocalSize */
int rdyb_vec_get_block_size(Vec v, int32_t *bs);    /* VecGetBlockSize */
int rdyb_vec_get_state(Vec v, int64_t *state);      /* PetscObjectStateGet: changes whenever the array was handed out for writing */
int rdyb_vec_zero(Vec v);                           /* VecZeroEntries */
int rdyb_vec_copy(Vec x, Vec y);                    /* VecCopy(x, y) */
#endif

/* ---- PetscOperator (include/private/rdyoperatorimpl.h:45-90) ---- */
#define RDYB_MAX_NAME_LEN 128 /* MAX_NAME_LEN, include/private/rdycoreimpl.h */

typedef struct {
  char name[RDYB_MAX_NAME_LEN + 1];
  Vec  vec;
} RdybPetscOperatorField;

typedef struct {
  PetscInt                num_fields, capacity;
  RdybPetscOperatorField *fields;
} RdybPetscOperatorFields;

typedef struct RdybPetscOperator_ *RdybPetscOperator;
struct RdybPetscOperator_ {
  void                   *context;
  RdybPetscOperatorFields fields;
  PetscBool               is_composite;
  PetscErrorCode (*apply)(void *, RdybPetscOperatorFields, PetscReal, Vec, Vec);
  PetscErrorCode (*destroy)(void *);
};

/* src/petsc_operator.c:11-21, 25-35, 42-47, 75-98, 105-115, 130-166 */
int rdyb_petscop_create(void *context, PetscErrorCode (*apply)(void *, RdybPetscOperatorFields, PetscReal, Vec, Vec),
                        PetscErrorCode (*destroy)(void *), RdybPetscOperator *op);
int rdyb_petscop_destroy(RdybPetscOperator *op);
int rdyb_petscop_apply(RdybPetscOperator op, PetscReal dt, Vec u_local, Vec f_global);
int rdyb_petscop_set_field(RdybPetscOperator op, const char *name, Vec vec); /* the operator does NOT own the Vec */
int rdyb_petscop_create_composite(RdybPetscOperator *op);
int rdyb_petscop_composite_add_sub(RdybPetscOperator op, RdybPetscOperator sub_op); /* the composite takes ownership */
int rdyb_petscop_fields_get(RdybPetscOperatorFields fields, const char *name, Vec *vec); /* *vec = NULL if absent */

#ifndef RDYB_HAVE_RDYCORE
/* outside an RDycore build the reference's own names read straight through */
#define PetscOperator RdybPetscOperator
#define PetscOperatorFields RdybPetscOperatorFields
#define PetscOperatorCreate rdyb_petscop_create
#define PetscOperatorDestroy rdyb_petscop_destroy
#define PetscOperatorApply rdyb_petscop_apply
#define PetscOperatorSetField rdyb_petscop_set_field
#define PetscOperatorCreateComposite rdyb_petscop_create_composite
#define PetscOperatorCompositeAddSub rdyb_petscop_composite_add_sub
#define PetscOperatorFieldsGet rdyb_petscop_fields_get
#endif

/* ---- factories ----
 * rdyb_create_flux_operator replaces CreatePetscFluxOperator and CreatePetscFluxHROperator (config->well_balancing picks):
 * a composite holding ONE sub-operator whose apply runs the fused CUDA kernel (interior + every boundary + sources) on
 * `device`.  boundary_values / boundary_fluxes / boundary_fluxes_accum are the Operator's per-boundary sequential Vecs
 * (src/operator.c:368-379; block size 3, one block per boundary edge); they stay owned by the caller.  `diagnostics` is
 * the Operator's CourantNumberDiagnostics (running max, src/swe/swe_petsc.c:289-296); may be NULL.
 * The views are read at create time only (nothing is retained).
 *
 * rdyb_create_source_operator replaces CreatePetscSourceOperator / CreatePetscSourceHROperator: external_sources
 * (block size 3) and material_properties (block size 1: Manning's n) are the Operator's Vecs in owned order
 * (src/operator.c:93-96); they are re-read whenever their state counter changed since the last apply.
 * It must be given the flux operator it pairs with: the two share one GPU operator.
 *
 * Split at the ABI, fused underneath.  With f_global zeroed by the caller (VecZeroEntries(F), src/rdysetup.c:1130):
 *   flux.apply   : uploads u_local, runs the fused kernel once, leaves the FLUX DIVERGENCE in f_global (so that the
 *                  reference's VecCopy(F, flux_divergence), src/operator.c:662, sees what it expects), fills
 *                  boundary_fluxes / boundary_fluxes_accum, updates the Courant diagnostics;
 *   source.apply : stores the complete right-hand side (flux divergence + sources) of that same kernel run in f_global
 *                  and fills the "primitive_variables" field when it is set.  No second kernel runs.
 * rdyb_petscop_set_fused(flux_op, 1) skips the intermediate: flux.apply writes the complete right-hand side and
 * source.apply does nothing (one device->host copy fewer per stage). */
int rdyb_create_flux_operator(const RdybConfig *config, const RdybMeshView *mesh, int32_t num_boundaries, const RdybBoundaryView *boundaries,
                              Vec *boundary_values, Vec *boundary_fluxes, Vec *boundary_fluxes_accum, RdybCourantDiagnostics *diagnostics,
                              int device, RdybPetscOperator *flux_op);
int rdyb_create_source_operator(const RdybConfig *config, const RdybMeshView *mesh, Vec external_sources, Vec material_properties,
                                RdybPetscOperator flux_op, RdybPetscOperator *source_op);
int rdyb_petscop_set_fused(RdybPetscOperator flux_op, int fused);
/* the GPU operator behind a flux (or source) operator made by the factories above: for kernel selection, device-resident
 * stepping, halo setup.  Borrowed. */
int rdyb_petscop_get_device_operator(RdybPetscOperator op, RdybOperator *device_op);

#ifdef __cplusplus
}
#endif
#endif /* RDYB200_PETSC_H */
