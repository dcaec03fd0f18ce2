/* Minimal stand-in for <petsc.h>: TEST INFRASTRUCTURE ONLY.
 * Lets the reference's own swe_petsc.c / petsc_operator.c /
 * operator_{fluxes,sources}_petsc.c compile, from where they lie under
 * /root/reference, without PETSc, libCEED or MPI.  Vec is a plain host array
 * with a block size; everything PETSc would do in serial is done literally. */
#ifndef RDYB_REF_SHIM_PETSC_H
#define RDYB_REF_SHIM_PETSC_H
#include <math.h>
#include <stddef.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include <stdarg.h>

typedef double PetscReal;
typedef double PetscScalar;
typedef int    PetscInt;
typedef int    PetscMPIInt;
typedef int    PetscErrorCode;
typedef int    PetscClassId;
typedef int    PetscLogEvent;
typedef enum { PETSC_FALSE = 0, PETSC_TRUE = 1 } PetscBool;
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
#define PetscInt_FMT "d"
#define PETSC_SUCCESS 0
#define PETSC_ERR_USER 83
#define PETSC_ERR_ARG_SIZ 60
#define PETSC_COMM_WORLD 1
#define PETSC_COMM_SELF 2
#define PETSC_VIEWER_STDOUT_WORLD NULL
#define PETSC_MAX_PATH_LEN 4096
#define PETSC_EXTERN extern
#define PETSC_INTERN extern

/* opaque handles that only appear in struct members / prototypes */
typedef struct _shim_DM          *DM;
typedef struct _shim_TS          *TS;
typedef struct _shim_Mat         *Mat;
typedef struct _shim_IS          *IS;
typedef struct _shim_SF          *PetscSF;
typedef struct _shim_Section     *PetscSection;
typedef struct _shim_Viewer      *PetscViewer;
typedef struct _shim_Bag         *PetscBag;
typedef struct _shim_Object      *PetscObject;
typedef struct _shim_VAF         *PetscViewerAndFormat;
typedef int                       PetscViewerFormat;
typedef int                       PetscFileMode;
typedef int                       InsertMode;
typedef const char               *VecType;
typedef struct _shim_VecScatter  *VecScatter;
typedef struct _shim_ISL2G       *ISLocalToGlobalMapping;
typedef struct _shim_Options     *PetscOptions;
#define PETSCHEADER(ops) int shim_header_
#define ADD_VALUES 2
#define INSERT_VALUES 1

/* Vec: a host array */
struct _shim_Vec {
  PetscScalar *a;
  PetscInt     n, bs;
  DM           dm;
};
typedef struct _shim_Vec *Vec;

#define PetscFunctionBegin
#define PetscFunctionBeginUser
#define PetscFunctionReturn(x) return (x)
#define PetscCall(x)                         \
  do {                                       \
    PetscErrorCode ierr_shim_ = (x);         \
    if (ierr_shim_) return ierr_shim_;       \
  } while (0)
#define PetscCallMPI(x) PetscCall(x)
#define PetscCheck(cond, comm, err, ...)     \
  do {                                       \
    if (!(cond)) {                           \
      fprintf(stderr, __VA_ARGS__);          \
      fprintf(stderr, "\n");                 \
      return (err);                          \
    }                                        \
  } while (0)
#define PetscAssert(cond, comm, err, ...) PetscCheck(cond, comm, err, __VA_ARGS__)
#define PetscCalloc1(n, p) ((*(p) = calloc((size_t)((n) > 0 ? (n) : 1), sizeof(**(p)))) ? 0 : 55)
#define PetscMalloc1(n, p) PetscCalloc1(n, p)
#define PetscFree(p) (free(p), (p) = NULL, 0)
#define PetscRealloc(sz, p) ((*(p) = realloc(*(p), (sz))) ? 0 : 55)
#define PetscPowReal(a, b) pow(a, b)
#define PetscSqrtReal(a) sqrt(a)
#define PetscAbsReal(a) fabs(a)
#define PetscAbsScalar(a) fabs(a)
#define PetscMin(a, b) (((a) < (b)) ? (a) : (b))
#define PetscMax(a, b) (((a) < (b)) ? (b) : (a))
#define PetscObjectGetComm(obj, comm) (*(comm) = PETSC_COMM_WORLD, 0)
#define PetscObjectComm(obj) PETSC_COMM_WORLD
#define PetscFPrintf(comm, f, ...) (fprintf(f, __VA_ARGS__), 0)
#define PetscPrintf(comm, ...) (printf(__VA_ARGS__), 0)

static inline PetscErrorCode PetscOptionsGetBool(void *a, void *b, const char *n, PetscBool *v, PetscBool *set) {
  (void)a; (void)b; (void)n; (void)v;
  if (set) *set = PETSC_FALSE;
  return 0;
}
static inline PetscErrorCode VecGetArray(Vec v, PetscScalar **a) { *a = v->a; return 0; }
static inline PetscErrorCode VecRestoreArray(Vec v, PetscScalar **a) { (void)v; *a = NULL; return 0; }
static inline PetscErrorCode VecGetBlockSize(Vec v, PetscInt *bs) { *bs = v->bs; return 0; }
static inline PetscErrorCode VecGetSize(Vec v, PetscInt *n) { *n = v->n; return 0; }
static inline PetscErrorCode VecGetLocalSize(Vec v, PetscInt *n) { *n = v->n; return 0; }
static inline PetscErrorCode VecGetDM(Vec v, DM *dm) { *dm = v->dm; return 0; }
static inline PetscErrorCode VecZeroEntries(Vec v) { memset(v->a, 0, sizeof(PetscScalar) * (size_t)v->n); return 0; }
static inline PetscErrorCode VecCopy(Vec x, Vec y) { memcpy(y->a, x->a, sizeof(PetscScalar) * (size_t)x->n); return 0; }
/* y = alpha x + y, as PETSc's serial kernel does (BLAS daxpy semantics) */
static inline PetscErrorCode VecAXPY(Vec y, PetscScalar alpha, Vec x) {
  for (PetscInt i = 0; i < y->n; ++i) y->a[i] += alpha * x->a[i];
  return 0;
}
static inline PetscErrorCode VecView(Vec v, void *viewer) { (void)v; (void)viewer; return 0; }
/* Serial DM stand-in for the second-order path (ApplyInteriorFlux2R, src/swe/swe_petsc.c:170-210): a local vector of
 * n_local blocks, and DMLocalToGlobal(ADD_VALUES) adding the owned rows of a local vector to the global one (the ghost rows
 * would travel to their owners; a serial mesh has none). */
struct _shim_DM {
  PetscInt         n_local, bs;
  const PetscBool *is_owned;
  const PetscInt  *local_to_owned;
};
static inline PetscErrorCode DMGetLocalVector(DM dm, Vec *v) {
  *v = calloc(1, sizeof(**v));
  (*v)->n = dm->n_local * dm->bs; (*v)->bs = dm->bs; (*v)->dm = dm;
  (*v)->a = calloc((size_t)((*v)->n > 0 ? (*v)->n : 1), sizeof(PetscScalar));
  return 0;
}
static inline PetscErrorCode DMRestoreLocalVector(DM dm, Vec *v) { (void)dm; free((*v)->a); free(*v); *v = NULL; return 0; }
static inline PetscErrorCode DMLocalToGlobal(DM dm, Vec l, InsertMode m, Vec g) {
  for (PetscInt c = 0; c < dm->n_local; ++c) {
    if (!dm->is_owned[c]) continue;
    for (PetscInt k = 0; k < dm->bs; ++k) {
      if (m == ADD_VALUES) g->a[dm->bs * dm->local_to_owned[c] + k] += l->a[dm->bs * c + k];
      else g->a[dm->bs * dm->local_to_owned[c] + k] = l->a[dm->bs * c + k];
    }
  }
  return 0;
}
#endif
