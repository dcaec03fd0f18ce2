/* swe_oracle.h — CPU oracle for the SWE right-hand-side path.  TEST INFRASTRUCTURE ONLY.
 *
 * Two shared libraries export this same interface:
 *   oracle/_ref/libswe_oracle_ref.so   the reference's own src/swe/swe_petsc.c, src/swe/swe_roe_flux_petsc.h,
 *                                      src/petsc_operator.c, src/operator_{fluxes,sources}_petsc.c compiled where
 *                                      they lie under /root/reference against oracle/ref_shim (a ~150-line PETSc
 *                                      stand-in), driven by oracle/ref_driver.c        -> kind "reference"
 *   oracle/libswe_oracle_port.so       oracle/swe_oracle.c, a from-scratch restatement -> kind "port"
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load them.
 * The product (librdyb200.so) never links, loads or calls anything in oracle/.
 *
 * Layouts are the reference's: AoS [h,hu,hv] Vecs with block size 3; u_local in the mesh view's local
 * cell order, everything "global" in owned order (cells.local_to_owned).
 */
#ifndef SWE_ORACLE_H
#define SWE_ORACLE_H
#include "rdyb200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct SweOracle_ *SweOracle;

const char *swe_oracle_kind(void);
/* CreateOperator (src/operator.c:348-417) for the PETSc backend on a mesh view */
int  swe_oracle_create(const RdybConfig *config, const RdybMeshView *mesh, int32_t num_boundaries, const RdybBoundaryView *boundaries,
                       SweOracle *oracle);
void swe_oracle_destroy(SweOracle oracle);
/* borrowed pointer to an internal Vec's array.  names: "u_global","u_local","rhs","external_sources",
 * "material_properties","flux_divergence","primitive_variables","src_inst"; per boundary b:
 * "boundary_values","boundary_fluxes","boundary_fluxes_accum" */
double *swe_oracle_array(SweOracle oracle, const char *name, int32_t boundary, int64_t *length);
/* OperatorRHSFunction (src/rdysetup.c:1120-1139): rhs <- 0; u_local[owned] <- u_global (the serial part of
 * DMGlobalToLocal; ghost entries of u_local are the caller's job); reset diagnostics; ApplyOperator */
int swe_oracle_rhs(SweOracle oracle, double dt);
/* TSEULER stage update: u_global <- u_global + dt*rhs (PETSc TSStep_Euler: VecAYPX(update, dt, U); VecCopy) */
int swe_oracle_euler_update(SweOracle oracle, double dt);
/* serial convenience: nsteps x (rhs + euler_update) */
int swe_oracle_step_euler(SweOracle oracle, double dt, int32_t nsteps);
/* serial classical RK4 (PETSc TSRK4 tableau), for the "next" row */
int swe_oracle_step_rk4(SweOracle oracle, double dt, int32_t nsteps);
/* max Courant number since the last rhs call began (diagnostics reset there) */
int swe_oracle_courant(SweOracle oracle, RdybCourantDiagnostics *diags);

#ifdef __cplusplus
}
#endif
#endif
