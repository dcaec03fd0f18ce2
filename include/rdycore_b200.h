/* rdycore_b200.h — RDycore's public lifecycle (include/rdycore.h) on top of the B200 SWE path.
 *
 * Same names, argument meaning and error behaviour as include/rdycore.h for the calls the hot path needs:
 *   RDyInit / RDyFinalize / RDyInitialized          include/rdycore.h:20-24
 *   RDyCreate / RDySetup / RDyDestroy               :30-32   (src/rdycore.c, src/rdysetup.c:1373-1575)
 *   RDyAdvance / RDyFinished                        :51-55   (src/rdyadvance.c:261-400)
 *   time queries                                    :60-69   (src/rdyadvance.c:402-456)
 *   RDyGetNum*, RDyGetOwnedCell*                    :71-84   (src/rdydata.c)
 *   RDySetFlowDirichletBoundaryValues               :95      (src/rdydata.c:88-106)
 *   RDySet{Regional,Domain}{Water,XMomentum,YMomentum}Source, RDySet{Regional,Domain}ManningsN  :104-118
 *   RDySetInitialConditions                         :119     (src/rdydata.c)
 * The state lives on the GPU between calls; RDyAdvance runs every stage of a coupling interval on the device (fused
 * Euler kernel or RHS + RK4 stage kernels) and only the adaptive-dt Courant number crosses back to the host.
 *
 * What is NOT here, on purpose: the YAML deck machinery beyond the scalar sections below, mesh files,
 * regions/materials/flow_conditions lists, output, checkpoints, time series, sediment, AMR, ensembles.  RDyCreate
 * reads the physics / numerics / time sections of an RDycore deck (a strict subset of the reference's YAML schema,
 * src/yaml_input.c:40-190: nested maps with scalar leaves; list sections are skipped); the mesh, boundaries, Manning's n
 * and initial state arrive through calls: RDyB200SetMesh before RDySetup, the rdycore.h setters after it.
 * MPI_Comm is an int here (no MPI in this image); pass 0.  One RDy per process per GPU.
 */
#ifndef RDYCORE_B200_H
#define RDYCORE_B200_H

#include "rdyb200_mesh.h"
#include "rdyb200_petsc.h"

#ifdef __cplusplus
extern "C" {
#endif

#ifndef RDYB_HAVE_PETSC
typedef int MPI_Comm;
#endif

typedef struct _p_RDy *RDy;

/* include/rdycore.h:58 */
typedef enum { RDY_TIME_UNSET = 0, RDY_TIME_SECONDS, RDY_TIME_MINUTES, RDY_TIME_HOURS, RDY_TIME_DAYS, RDY_TIME_MONTHS, RDY_TIME_YEARS } RDyTimeUnit;

PetscErrorCode RDyInit(int argc, char *argv[], const char *help);
PetscInt       RDyFinalize(void);
PetscBool      RDyInitialized(void);

PetscErrorCode RDyCreate(MPI_Comm comm, const char *config_file, RDy *rdy);
PetscErrorCode RDySetup(RDy rdy);
PetscErrorCode RDyDestroy(RDy *rdy);

PetscErrorCode RDyAdvance(RDy rdy);
PetscBool      RDyFinished(RDy rdy);

PetscErrorCode RDyGetTimeUnit(RDy rdy, RDyTimeUnit *unit);
PetscErrorCode RDyGetTime(RDy rdy, RDyTimeUnit unit, PetscReal *time);
PetscErrorCode RDyGetTimeStep(RDy rdy, RDyTimeUnit unit, PetscReal *dt);
PetscErrorCode RDyConvertTime(RDyTimeUnit unit_from, PetscReal t_from, RDyTimeUnit unit_to, PetscReal *t_to);
PetscErrorCode RDyGetStep(RDy rdy, PetscInt *step);
PetscErrorCode RDyGetCouplingInterval(RDy rdy, RDyTimeUnit unit, PetscReal *interval);
PetscErrorCode RDySetCouplingInterval(RDy rdy, RDyTimeUnit unit, PetscReal interval);

PetscErrorCode RDyGetNumGlobalCells(RDy rdy, PetscInt *n);
PetscErrorCode RDyGetNumOwnedCells(RDy rdy, PetscInt *n);
PetscErrorCode RDyGetNumBoundaryConditions(RDy rdy, PetscInt *n);
PetscErrorCode RDyGetNumBoundaryEdges(RDy rdy, const PetscInt boundary_index, PetscInt *n);
PetscErrorCode RDyGetBoundaryConditionFlowType(RDy rdy, const PetscInt boundary_index, PetscInt *bc_type);
PetscErrorCode RDyGetBoundaryID(RDy rdy, const PetscInt boundary_index, PetscInt *id);

PetscErrorCode RDyGetOwnedCellHeights(RDy rdy, const PetscInt size, PetscReal *values);
PetscErrorCode RDyGetOwnedCellXMomenta(RDy rdy, const PetscInt size, PetscReal *values);
PetscErrorCode RDyGetOwnedCellYMomenta(RDy rdy, const PetscInt size, PetscReal *values);
PetscErrorCode RDyGetOwnedCellXCentroids(RDy rdy, const PetscInt size, PetscReal *values);
PetscErrorCode RDyGetOwnedCellYCentroids(RDy rdy, const PetscInt size, PetscReal *values);
PetscErrorCode RDyGetOwnedCellAreas(RDy rdy, const PetscInt size, PetscReal *values);

PetscErrorCode RDySetFlowDirichletBoundaryValues(RDy rdy, const PetscInt boundary_index, const PetscInt num_edges, const PetscInt ndof,
                                                 PetscReal *values);
PetscErrorCode RDySetDomainWaterSource(RDy rdy, PetscInt size, PetscReal *values);
PetscErrorCode RDySetDomainXMomentumSource(RDy rdy, PetscInt size, PetscReal *values);
PetscErrorCode RDySetDomainYMomentumSource(RDy rdy, PetscInt size, PetscReal *values);
PetscErrorCode RDySetRegionalWaterSource(RDy rdy, const PetscInt region_idx, PetscInt size, PetscReal *values);
PetscErrorCode RDySetRegionalXMomentumSource(RDy rdy, const PetscInt region_idx, PetscInt size, PetscReal *values);
PetscErrorCode RDySetRegionalYMomentumSource(RDy rdy, const PetscInt region_idx, PetscInt size, PetscReal *values);
PetscErrorCode RDySetDomainManningsN(RDy rdy, const PetscInt size, PetscReal *values);
PetscErrorCode RDySetRegionalManningsN(RDy rdy, const PetscInt region_idx, const PetscInt size, PetscReal *values);
/* ic: block size 3, owned order (the prognostic Vec of RDyCreatePrognosticVec) */
PetscErrorCode RDySetInitialConditions(RDy rdy, Vec ic);
PetscErrorCode RDyCreatePrognosticVec(RDy rdy, Vec *prog_vec);

/* ---- what replaces DMPlex + the list sections of the deck ---- */
/* The mesh and its boundaries (borrowed until RDyDestroy; call before RDySetup).  `halo` is the rank's halo plan for a
 * partitioned mesh or NULL.  Boundary edges of the mesh named by no boundary get the reference's auto-generated reflecting
 * boundary (src/rdysetup.c:425-441, 768-773) — the caller's list must already contain it (rdyb_hostmesh helpers do). */
PetscErrorCode RDyB200SetMesh(RDy rdy, RdybHostMesh mesh, PetscInt num_boundaries, const RdybBoundaryView *boundaries);
/* regions: owned-cell id lists (RDyRegion.owned_cell_global_ids, include/private/rdyregionimpl.h:8-25); copied */
PetscErrorCode RDyB200AddRegion(RDy rdy, PetscInt num_owned_cells, const PetscInt *owned_cell_ids, PetscInt *region_idx);
/* CUDA device ordinal (default 0) and kernel choice (rdyb_operator_set_kernel; default K4 with 224-cell tiles) */
PetscErrorCode RDyB200SetDevice(RDy rdy, int device);
PetscErrorCode RDyB200SetKernel(RDy rdy, PetscInt kind, PetscInt tile_cells);
/* the operator behind the RDy (valid after RDySetup): for halo setup, diagnostics, device pointers */
PetscErrorCode RDyB200GetOperator(RDy rdy, RdybOperator *op);
/* Checkpoints, binary format (include/rdyb200_io.h).  Write: WriteCheckpointFile (src/checkpoint.c:153-212) to "<filename_base>.bin" —
 * metadata {nproc, t, dt, step} and the solution in natural (global cell id) order; on a partitioned mesh every process calls it with its
 * rank and the global cell count, and writes its owned rows in place (rank 0 also writes the headers).  Read: ReadCheckpointFile +
 * ConsumeMetadata (:80-97, 258-320): state, dt and — unless `reinitialize`, the deck's restart.reinitialize — time and step. */
PetscErrorCode RDyB200WriteCheckpoint(RDy rdy, const char *filename_base, PetscInt nproc, PetscInt rank, PetscInt num_global_cells);
PetscErrorCode RDyB200ReadCheckpoint(RDy rdy, const char *filename, PetscBool reinitialize);
/* max Courant number of the last right-hand-side evaluation (GetOperatorDiagnostics, src/operator.c:889-893) */
PetscErrorCode RDyB200GetMaxCourantNumber(RDy rdy, PetscReal *courant);

#ifdef __cplusplus
}
#endif
#endif /* RDYCORE_B200_H */
