/* rdyb200.h — C-ABI of the B200-native shallow-water (SWE) right-hand-side path.
 *
 * Drop-in boundary for RDycore's PETSc operator backend.  Every entry point
 * is `extern "C"`, takes plain pointers and sizes, returns an int error code
 * (0 = success, like PetscErrorCode) and never exits the process.  The last
 * error text of the calling thread is available from rdyb_last_error().
 *
 * Citations are relative to the RDycore source tree (reference):
 *   Operator / PetscOperator ABI      include/private/rdyoperatorimpl.h:75-90,107-213
 *   flux factories                    src/operator_fluxes_petsc.c:17-95
 *   source factories                  src/operator_sources_petsc.c:13-45
 *   CPU kernels replaced              src/swe/swe_petsc.c, src/swe/swe_roe_flux_petsc.h
 *   RHS callback + TS stage update    src/rdysetup.c:1120-1172, 1183-1196
 *   mesh quantities consumed          include/private/rdymeshimpl.h:26-75,106-140
 *   data access                       src/operator.c:1045-1086,1203-1290,1394-1464
 *   diagnostics                       src/operator.c:757-768,867-893
 */
#ifndef RDYB200_H
#define RDYB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- error codes (values of the PETSc codes the reference raises) ---- */
#define RDYB_SUCCESS 0
#define RDYB_ERR_MEM 55      /* PETSC_ERR_MEM */
#define RDYB_ERR_SUP 56      /* PETSC_ERR_SUP: unsupported configuration */
#define RDYB_ERR_ARG_SIZ 60  /* PETSC_ERR_ARG_SIZ */
#define RDYB_ERR_ARG 62      /* PETSC_ERR_ARG_WRONG */
#define RDYB_ERR_USER 83     /* PETSC_ERR_USER: what PetscCheck(..., PETSC_ERR_USER, ...) raises */
#define RDYB_ERR_CUDA 97     /* PETSC_ERR_GPU */
#define RDYB_ERR_NCCL 98

/* ---- enums: same numeric values as the reference ---- */
/* include/rdycore.h:133-139 (RDyConditionType) */
enum { RDYB_CONDITION_DIRICHLET = 0, RDYB_CONDITION_NEUMANN = 1, RDYB_CONDITION_REFLECTING = 2, RDYB_CONDITION_CRITICAL_OUTFLOW = 3 };
/* include/private/rdyconfigimpl.h:52-56 (RDyFlowSourceMethod) */
enum { RDYB_SOURCE_SEMI_IMPLICIT = 0, RDYB_SOURCE_IMPLICIT_XQ2018 = 1 };
/* include/private/rdyconfigimpl.h:58-62 (RDyWellBalanceMethod); BS2002 falls back to NONE in the PETSc backend (src/operator.c:202-209) */
enum { RDYB_WELL_BALANCING_NONE = 0, RDYB_WELL_BALANCING_BS2002 = 1, RDYB_WELL_BALANCING_HR = 2 };
/* include/private/rdyconfigimpl.h:110-115 (RDyNumericsTemporal) */
enum { RDYB_TEMPORAL_EULER = 0, RDYB_TEMPORAL_RK4 = 1 };
/* config.numerics.second_order + limiter (include/private/rdyconfigimpl.h; resolved as in src/swe/swe_petsc.c:357-369):
 * first order, or MUSCL reconstruction with the minmod (default), no, or van Leer limiter */
enum { RDYB_RECON_FIRST_ORDER = 0, RDYB_RECON_MUSCL_MINMOD = 1, RDYB_RECON_MUSCL_NONE = 2, RDYB_RECON_MUSCL_VANLEER = 3 };

/* Physics / numerics parameters the kernels read.  Filled from RDyConfig:
 * config.physics.flow.{tiny_h,h_anuga_regular,source.method,source.xq2018_threshold,well_balancing},
 * config.numerics.temporal.  Defaults: src/yaml_input.c:854-865. */
typedef struct {
  double  tiny_h;           /* default 1e-7 */
  double  h_anuga_regular;  /* default 0.0 */
  double  xq2018_threshold; /* default 1e-10 */
  int32_t source_method;    /* RDYB_SOURCE_* */
  int32_t well_balancing;   /* RDYB_WELL_BALANCING_* */
  int32_t temporal;         /* RDYB_TEMPORAL_* */
  int32_t reconstruction;   /* RDYB_RECON_* (0 = first order) */
} RdybConfig;

/* A borrowed, read-only view of the RDyMesh arrays the path consumes
 * (include/private/rdymeshimpl.h).  Local numbering, PetscInt = int32.
 * A maintainer fills it with pointer assignments from RDyMesh; nothing is
 * retained after rdyb_operator_create() returns. */
typedef struct {
  int32_t num_cells;           /* RDyMesh.num_cells (owned + ghost) */
  int32_t num_owned_cells;     /* RDyMesh.num_owned_cells */
  int32_t num_edges;           /* RDyMesh.num_edges */
  int32_t num_internal_edges;  /* RDyMesh.num_internal_edges */
  /* cells, length num_cells */
  const int32_t *cell_is_owned;       /* RDyCells.is_owned */
  const int32_t *cell_local_to_owned; /* RDyCells.local_to_owned */
  const int64_t *cell_global_ids;     /* RDyCells.global_ids (may be NULL => local id) */
  const double  *cell_areas;          /* RDyCells.areas */
  const double  *cell_dz_dx;          /* RDyCells.dz_dx */
  const double  *cell_dz_dy;          /* RDyCells.dz_dy */
  const double  *cell_zc;             /* HR bed elevation, swe_petsc.c:1208-1224 (NULL unless HR) */
  const double  *cell_xc, *cell_yc;   /* RDyCells.centroids[].X[0..1]; used only to order cells for locality (may be NULL) */
  /* edges, length num_edges unless noted */
  const int32_t *edge_cell_ids;       /* RDyEdges.cell_ids, [2e] left, [2e+1] right or -1 */
  const double  *edge_lengths;        /* RDyEdges.lengths */
  const double  *edge_cn, *edge_sn;   /* RDyEdges.cn, sn */
  const int32_t *internal_edge_ids;   /* RDyEdges.internal_edge_ids, length num_internal_edges */
  const int64_t *edge_global_ids;     /* RDyEdges.global_ids (may be NULL => local id) */
  /* second order only (NULL otherwise): edge midpoints 0.5*(v0+v1) as ReconstructFaceValues forms them from
   * RDyEdges.vertex_ids / RDyVertices.points (src/operator_fluxes_ceed.c:1162-1166); cell_xc/cell_yc are then required too */
  const double  *edge_xm, *edge_ym;
} RdybMeshView;

/* One RDyBoundary (include/private/rdyboundaryimpl.h:9-16) with the flow
 * type of its RDyCondition (include/private/rdyconditionimpl.h:12-24). */
typedef struct {
  int32_t        id, index;
  int32_t        num_edges;
  const int32_t *edge_ids; /* local edge ids */
  int32_t        bc_type;  /* RDYB_CONDITION_* */
} RdybBoundaryView;

/* CourantNumberDiagnostics, include/private/rdyoperatorimpl.h:21-25 */
typedef struct {
  double  max_courant_num;
  int64_t global_edge_id;
  int64_t global_cell_id;
} RdybCourantDiagnostics;

/* Halo plan for one rank: which owned cells go to which peer and which ghost
 * cells arrive from which peer (cell ids are LOCAL ids of the RdybMeshView).
 * Replaces the PetscSF behind DMGlobalToLocal (src/rdysetup.c:1133-1134). */
typedef struct {
  int32_t        num_peers;
  const int32_t *peer_ranks;   /* [num_peers] */
  const int32_t *send_offsets; /* [num_peers+1] into send_cells */
  const int32_t *send_cells;   /* local ids of owned cells, per peer, ascending global id */
  const int32_t *recv_offsets; /* [num_peers+1] into recv_cells */
  const int32_t *recv_cells;   /* local ids of ghost cells, per peer, ascending global id */
} RdybHaloView;

/* Vertex data for Bradford-Sanders well balancing (rdyb_operator_enable_bs2002): RDyVertices.points[].X[2],
 * RDyCells.vertex_offsets / vertex_ids, RDyEdges.vertex_ids (include/private/rdymeshimpl.h:40-44,96,117), view-local numbering */
typedef struct {
  int32_t        num_vertices;
  const double  *vertex_z;            /* [num_vertices] */
  const int32_t *cell_vertex_offsets; /* [num_cells+1] */
  const int32_t *cell_vertex_ids;
  const int32_t *edge_vertex_ids;     /* [2*num_edges]: beginning and end vertex of every edge */
} RdybVertexView;

typedef struct RdybOperator_ *RdybOperator;

const char *rdyb_last_error(void);
int         rdyb_version(int *major, int *minor);
/* number of kernels this library launched on the calling process since load */
int64_t rdyb_kernel_launch_count(void);

/* ---- construction: CreateOperator (src/operator.c:348-417) for the SWE/PETSc case ----
 * `device` = CUDA ordinal.  `cuda_stream` = a cudaStream_t to enqueue on (NULL => a stream the operator owns). */
int rdyb_operator_create(const RdybConfig *config, const RdybMeshView *mesh, int32_t num_boundaries, const RdybBoundaryView *boundaries,
                         const RdybHaloView *halo /* NULL for serial */, int device, void *cuda_stream, RdybOperator *op);
int rdyb_operator_destroy(RdybOperator *op);
int rdyb_operator_num_owned_cells(RdybOperator op, int32_t *n);
int rdyb_operator_num_local_cells(RdybOperator op, int32_t *n);
/* toggles: Courant diagnostics every apply (reference always does, swe_petsc.c:289-296), flux-divergence /
 * primitive-variable outputs (src/operator.c:224-299), boundary flux accumulation (swe_petsc.c:623) */
int rdyb_operator_set_options(RdybOperator op, int courant, int flux_divergence, int primitive_variables, int boundary_fluxes);

/* kernel selection: kind 0 = K1 cell-centric gather (default), 1 = K2 shared-memory tiles of `tile_cells` cells,
 * 2 = K3 persistent bulk-async-staged tiles, 3 = K4 (K3 with compile-time tile shapes, tuned for tile_cells 224; shapes that do
 * not fit run K3).  With
 * config.reconstruction != 0 (second order) every kind >= 1 runs K2M, the tiled MUSCL kernel (tuned for tile_cells 128), kind 0
 * the cell-centric MUSCL gather; rdyb_operator_kernel_info reports which kernel the last launch ran. */
int rdyb_operator_set_kernel(RdybOperator op, int32_t kind, int32_t tile_cells);

/* at most `max_ctas` CTAs in the grid of the persistent tile kernels (0 = no cap; env RDYB_MAX_CTAS is the process-wide default).
 * Tests use it so that on a small mesh every CTA walks many tiles - all pipeline stages, both mbarrier parities - as in the bench. */
int rdyb_operator_set_max_ctas(RdybOperator op, int32_t max_ctas);

/* ---- operator data (host arrays in the reference's layouts) ---- */
/* SetOperatorBoundaryValues, src/operator.c:1045-1061: values[num_comp*e + c] -> components [comp_offset, comp_offset+num_comp) */
int rdyb_operator_set_boundary_values(RdybOperator op, int32_t boundary_index, int32_t comp_offset, int32_t num_comp, int32_t num_edges,
                                      const double *values);
/* ExtractOperatorBoundaryFluxes, src/operator.c:1069-1086: accumulated fluxes, values[c][e] flattened as c*num_edges+e */
int rdyb_operator_get_boundary_fluxes(RdybOperator op, int32_t boundary_index, int32_t num_edges, int accumulated, double *values);
/* Get/RestoreOperatorDomainExternalSource, src/operator.c:1394-1429: one component, owned order */
int rdyb_operator_set_external_source(RdybOperator op, int32_t component, int32_t num_owned, const double *values);
int rdyb_operator_get_external_source(RdybOperator op, int32_t component, int32_t num_owned, double *values);
/* the whole external_sources Vec at once: AoS values[3*i + c] in owned order (src/operator.c:93-96) */
int rdyb_operator_set_external_source_aos(RdybOperator op, int32_t num_owned, const double *values);
/* Get/RestoreOperatorRegionalExternalSource, src/operator.c:1203-1241: values[i] -> owned cell owned_cell_ids[i] */
int rdyb_operator_set_regional_external_source(RdybOperator op, int32_t component, int32_t n, const int32_t *owned_cell_ids, const double *values);
/* Get/RestoreOperatorDomainMaterialProperties (Manning's n), src/operator.c:1435-1464 */
int rdyb_operator_set_mannings(RdybOperator op, int32_t num_owned, const double *values);
int rdyb_operator_set_regional_mannings(RdybOperator op, int32_t n, const int32_t *owned_cell_ids, const double *values);

/* ---- state: the solution Vec (AoS [h,hu,hv], block size 3) ---- */
/* u_global: owned order, 3*num_owned doubles (rdy->u_global, src/rdydm.c:254-292) */
int rdyb_operator_set_state(RdybOperator op, const double *u_global);
int rdyb_operator_get_state(RdybOperator op, double *u_global);
/* RDyGetOwnedCell{Heights,XMomenta,YMomenta}, src/rdydata.c: one component, owned order */
int rdyb_operator_get_state_component(RdybOperator op, int32_t component, int32_t num_owned, double *values);

/* optional output Vecs of the last apply in the reference's AoS owned layout (3*num_owned doubles):
 * which = 0 flux_divergence ("riemannf", src/operator.c:224-260), 1 primitive_variables (:264-299), 2 F */
int rdyb_operator_get_field(RdybOperator op, int32_t which, double *values);
/* pinned host memory for host-Vec callers of rdyb_operator_apply_host */
int rdyb_host_alloc(size_t bytes, void **ptr);
int rdyb_host_free(void *ptr);

/* ---- the hot path ---- */
/* PetscOperator apply shape (include/private/rdyoperatorimpl.h:80) with HOST arrays:
 * u_local = AoS [num_cells][3] in the mesh view's local order (ghosts filled by the caller),
 * f_global = AoS [num_owned][3]; computes flux + source exactly as ApplyPetscOperator
 * (src/operator.c:656-672) after VecZeroEntries(F) (src/rdysetup.c:1130).  Copies in, runs, copies out. */
int rdyb_operator_apply_host(RdybOperator op, double dt, const double *u_local, double *f_global);
/* Same, on the device-resident state (ghosts refreshed by the halo exchange when a halo plan exists);
 * F stays on the device; f_global may be NULL or a host array to download into. */
int rdyb_operator_apply(RdybOperator op, double dt, double *f_global);
/* OperatorRHSFunction + TSEULER step fused, nsteps times: U <- U + dt*F(U).  Asynchronous; state stays resident. */
int rdyb_operator_step_euler(RdybOperator op, double dt, int32_t nsteps);
/* classical RK4 (TSRK4) built from the RHS-only kernel and AXPY stage kernels */
int rdyb_operator_step_rk4(RdybOperator op, double dt, int32_t nsteps);
int rdyb_operator_synchronize(RdybOperator op);

/* ---- Bradford-Sanders (2002) well balancing with the semantics of the reference's libCEED operator (the PETSc operator ignores the
 * option, src/operator.c:202-209, and so does rdyb_operator_create): a vertex pass (SWEEtaVertex, src/swe/swe_well_balancing.h:30)
 * before every RHS evaluation and g dhv^2 / 24 (cn, sn) on every edge flux (src/swe/swe_fluxes_ceed.h:15-19,
 * src/swe/swe_roe_flux_ceed.h:119-121).  Triangles, serial meshes, first order; runs on the cell-centric kernel K1. */
int rdyb_operator_enable_bs2002(RdybOperator op, const RdybVertexView *vertices);
int rdyb_operator_get_eta_vertices(RdybOperator op, int32_t num_vertices, double *eta);

/* ---- one coupling interval on the device: RDyAdvance without a host round trip (src/rdyadvance.c:261-383) ----
 * rdyb_operator_time_set puts (t, dt) into a device-resident control block.  rdyb_operator_advance_device then enqueues ONE CUDA
 * graph: a kernel that reduces the Courant slots of the previous interval's last step and applies the adaptive-dt rule
 * (:303-343; skipped when `adaptive` is 0 or no Courant number exists yet), and a device-side WHILE loop of fused Euler steps -
 * whole steps of dt, then one shortened step onto t + interval (TS_EXACTFINALTIME_MATCHSTEP) - whose kernels read dt from the
 * control block.  Nothing waits on the host; rdyb_operator_time_get (and every state getter) synchronises when asked.
 * Returns RDYB_ERR_SUP without doing anything when the operator has no device-side loop (partitioned mesh, second order, a
 * kernel other than K1 / K4): the caller then steps with rdyb_operator_step_euler and its own bookkeeping. */
int rdyb_operator_time_set(RdybOperator op, double t, double dt);
int rdyb_operator_time_get(RdybOperator op, double *t, double *dt, int64_t *steps, double *last_courant);
int rdyb_operator_advance_device(RdybOperator op, double interval, int adaptive, double target_courant, double max_increase_factor);
/* host waits for the device (stream / event synchronisations) this library performed since load: the "zero host syncs per
 * RDyAdvance" figure of the tests */
int64_t rdyb_host_sync_count(void);

/* ---- diagnostics: Reset/Update/GetOperatorDiagnostics, src/operator.c:757-768,867-893 ---- */
int rdyb_operator_reset_diagnostics(RdybOperator op);
int rdyb_operator_get_courant(RdybOperator op, RdybCourantDiagnostics *diags); /* local max since reset (syncs) */
/* UpdateOperatorDiagnostics, src/operator.c:867-885: max over the ranks of the NCCL communicator (collective); ids stay on the rank holding it */
int rdyb_operator_get_courant_global(RdybOperator op, RdybCourantDiagnostics *diags);
/* sum over owned cells of area*h (mass); deterministic two-stage reduction (syncs) */
int rdyb_operator_get_mass(RdybOperator op, double *mass);

/* Allocation-free mass sum.  NO COUNTERPART IN THE REFERENCE: added for callers that cannot
 * call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The reduction of
 * rdyb_operator_get_mass - sum_i area[i]*h[i] by 1024 fixed blocks and a fixed tree, the partial sums added in index order - with its
 * scratch (1024 doubles = 8 KiB) in a file-scope `static __device__` array of the library (swe_noalloc.cu) instead of the operator's
 * cudaMalloc'ed buffer, and the final sum done on the device.
 *   h, area      DEVICE pointers to num_cells doubles each, on the CURRENT device (e.g. plane 0 of rdyb_operator_device_pointer)
 *   num_cells    0 .. RDYB_NOALLOC_MAX_CELLS; beyond that RDYB_ERR_ARG_SIZ and nothing is launched
 *   mass         DEVICE pointer to one double, written by the second kernel
 *   cuda_stream  cudaStream_t (NULL = the default stream); two kernels are enqueued on it, the call does not synchronise
 * Calls no cudaMalloc / cudaFree / allocating library routine, copies nothing, never blocks.  Every call on a GPU shares the one static
 * array (each call overwrites all of it before reading it, so nothing has to be cleared): calls that overlap on the same GPU race -
 * order them on one stream or with events.  The first call on a device resolves the array's device address (cudaGetSymbolAddress,
 * cached per device): make that call before a stream capture begins. */
#define RDYB_NOALLOC_MAX_CELLS 1073741824 /* 2^30 cells */
int rdyb_mass_sum_noalloc(const double *h, const double *area, int64_t num_cells, double *mass, void *cuda_stream);

/* ---- multi-GPU: NCCL communicator for the halo exchange ---- */
int rdyb_nccl_unique_id(void *id128 /* 128 bytes out */);
int rdyb_operator_init_comm(RdybOperator op, int32_t nranks, int32_t rank, const void *id128);
/* run one halo exchange now (owned -> peers' ghosts) on the current state */
int rdyb_operator_halo_exchange(RdybOperator op);
/* Peer-memory halo for the fused Euler step (one process per GPU on one NVLink box): every rank exports a blob of
 * RDYB_HALO_BLOB_BYTES (CUDA IPC handles of its state buffers and flag words + its halo plan), the host gathers the blobs of
 * all ranks by any means (MPI_Allgather on the RDycore side) and hands the nranks*RDYB_HALO_BLOB_BYTES array to import.
 * Afterwards rdyb_operator_step_euler writes border cells straight into the neighbours' ghost slots over NVLink instead of
 * going through pack -> ncclSend/ncclRecv -> unpack.  Collective: all ranks must call step_euler with the same step counts. */
#define RDYB_HALO_BLOB_BYTES 512
int rdyb_operator_halo_export(RdybOperator op, int32_t rank, void *blob);
int rdyb_operator_halo_import(RdybOperator op, int32_t nranks, int32_t rank, const void *blobs);
/* clock64() ticks the in-kernel neighbour waits of this rank took since the last call (sum over the waiting CTAs, longest single
 * wait, number of waits = CTAs owning a frontier tile x launches); resets the counters.  The scaling diagnostic of bench.py: how much of a step is spent waiting for border cells. */
int rdyb_operator_get_wait_cycles(RdybOperator op, uint64_t *sum, uint64_t *longest, uint64_t *waits);

/* ---- raw device pointers (for zero-copy binding to VECCUDA / torch tensors) ---- */
/* which: 0..2 = h,hu,hv current state (length num_local), 3..5 = F (length num_owned) */
int rdyb_operator_device_pointer(RdybOperator op, int32_t which, void **ptr, int64_t *length);
/* timing helper: average kernel time of the last step_euler call's main kernel, measured with
 * CUDA events on the operator's stream (ms per launch) */
int rdyb_operator_last_kernel_ms(RdybOperator op, double *ms_per_launch, int32_t *launches);
/* text description of the kernel the last launch ran and of the device layout (a requested kernel falls back to a more
 * general one when the mesh does not fit its compiled tile shapes) */
int rdyb_operator_kernel_info(RdybOperator op, char *buf, int32_t len);

#ifdef __cplusplus
}
#endif
#endif /* RDYB200_H */
