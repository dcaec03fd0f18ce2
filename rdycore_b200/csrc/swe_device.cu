/* swe_device.cu — the thin C-ABI CUDA layer of the B200 SWE path (include/rdyb200.h).
 *
 * Owns device memory, streams, the NCCL communicator of the halo exchange and the kernel launches.  The host
 * logic above it (mesh, device layout, RDy lifecycle) is C.  There is no CPU fallback: every entry point that
 * computes needs a CUDA device and fails with RDYB_ERR_CUDA otherwise.
 */
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <atomic>
#include <algorithm>
#include <vector>

#include "device_layout.h"
#include "rdyb200.h"
#include "rdyb_error.h"
#include "swe_kernels.cuh"
#include "swe_tiled.cuh"
#include "swe_tiled_async.cuh"
#include "swe_tiled_fixed.cuh"
#include "swe_tiled_muscl.cuh"
#include "swe_tiled_muscl_async.cuh"
#include "swe_tiled_muscl_fixed.cuh"

static std::atomic<int64_t> g_launches{0};
static std::atomic<int64_t> g_host_syncs{0}; /* host waits for the device (stream / event synchronisations) since load */
static inline cudaError_t rdyb_stream_sync(cudaStream_t s) { g_host_syncs.fetch_add(1, std::memory_order_relaxed); return cudaStreamSynchronize(s); }
static inline cudaError_t rdyb_event_sync(cudaEvent_t e) { g_host_syncs.fetch_add(1, std::memory_order_relaxed); return cudaEventSynchronize(e); }

#define CU(call)                                                                                                   \
  do {                                                                                                             \
    cudaError_t e_ = (call);                                                                                       \
    if (e_ != cudaSuccess) return rdyb_fail(RDYB_ERR_CUDA, "CUDA error %s at %s:%d (%s)", cudaGetErrorName(e_), __FILE__, __LINE__, cudaGetErrorString(e_)); \
  } while (0)
#define LAUNCHED() (g_launches.fetch_add(1, std::memory_order_relaxed))

/* ---------------- NCCL, bound at run time so that the library loads on hosts without it ---------------- */
struct NcclApi {
  void *handle = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
static NcclApi g_nccl;

static int nccl_load() {
  if (g_nccl.handle) return RDYB_SUCCESS;
  const char *names[] = {"libnccl.so.2", "libnccl.so"};
  void *h = nullptr;
  for (const char *n : names) {
    h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (h) break;
  }
  if (!h) return rdyb_fail(RDYB_ERR_NCCL, "NCCL library not found (%s)", dlerror());
#define SYM(field, name)                                                                           \
  do {                                                                                             \
    *(void **)(&g_nccl.field) = dlsym(h, name);                                                    \
    if (!g_nccl.field) return rdyb_fail(RDYB_ERR_NCCL, "NCCL symbol %s not found", name);           \
  } while (0)
  SYM(GetUniqueId, "ncclGetUniqueId"); SYM(CommInitRank, "ncclCommInitRank"); SYM(CommDestroy, "ncclCommDestroy");
  SYM(GroupStart, "ncclGroupStart"); SYM(GroupEnd, "ncclGroupEnd"); SYM(Send, "ncclSend"); SYM(Recv, "ncclRecv");
  SYM(AllReduce, "ncclAllReduce"); SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
  g_nccl.handle = h;
  return RDYB_SUCCESS;
}
#define NC(call)                                                                                                   \
  do {                                                                                                             \
    ncclResult_t r_ = (call);                                                                                      \
    if (r_ != ncclSuccess) return rdyb_fail(RDYB_ERR_NCCL, "NCCL error at %s:%d (%s)", __FILE__, __LINE__, g_nccl.GetErrorString(r_)); \
  } while (0)

/* ---------------- the operator ---------------- */
struct RdybOperator_ {
  RdybConfig cfg;
  RdybLayout L;      /* host copy of the layout (index maps are needed by the setters) */
  RdybTiledLayout T; /* tiled layout (host copy of the sizes; arrays are uploaded) */
  int device = 0;
  cudaStream_t stream = nullptr, comm_stream = nullptr;
  bool own_stream = false;
  cudaEvent_t ev_pack = nullptr, ev_halo = nullptr, ev_t0 = nullptr, ev_t1 = nullptr;
  int32_t N = 0, NL = 0, NI = 0, stride = 0, width = 0, nbe = 0;
  /* device arrays */
  double *state[2] = {nullptr, nullptr}; /* [3][stride] each; `cur` is the solution */
  int cur = 0;
  double *F = nullptr, *src = nullptr, *mann = nullptr, *fdiv = nullptr, *prim = nullptr;
  double *inv_area = nullptr, *area = nullptr, *dzdx = nullptr, *dzdy = nullptr, *zc = nullptr;
  int32_t *ell_nbr = nullptr;
  double *ell_sn = nullptr, *ell_cn = nullptr, *ell_slen = nullptr;
  int32_t *btype = nullptr;
  double *bval = nullptr, *bflux = nullptr, *bacc = nullptr;
  int32_t *dev_to_local = nullptr, *dev_to_owned = nullptr, *send_cells = nullptr;
  double *stage = nullptr;    /* [3][stride] staging for host <-> device conversions */
  int32_t *stage_idx = nullptr;
  double *sendbuf = nullptr, *recvbuf = nullptr;
  double *cmax = nullptr;
  long long *carg = nullptr;
  int32_t ncblocks = 0, ncblocks_interior = 0;
  double *partial = nullptr;
  double *rk_k[4] = {nullptr, nullptr, nullptr, nullptr}, *rk_y = nullptr;
  /* second order (MUSCL): reconstruction kind (RDYB_RECON_*, 0 = first order), cell gradients [6][stride], per-slot geometry */
  int recon = 0;
  double *grad = nullptr, *m_gcx = nullptr, *m_gcy = nullptr, *m_dxs = nullptr, *m_dys = nullptr, *m_dxn = nullptr, *m_dyn = nullptr;
  double *t_edge_off = nullptr; /* [4][T.total_edges] face offsets per tile edge: the second-order path on the tiled layout (K2M) */
  int32_t *t_halo2 = nullptr;   /* two-ring staging of K4M (device_layout.h): second-ring cells per tile ... */
  uint16_t *t_halo_nbr = nullptr; /* ... and the staged position of every halo cell's neighbours; both NULL when the layout has none */
  double *t_halo_rec = nullptr;   /* [N][8] {1/area, LS weights}: one 64-byte record per cell for K4M's ring-1 gathers (swe_tiled_muscl_fixed.cuh) */
  bool courant_arg_tiled = false; /* the Courant arg of the last launch is a tile-edge index (else device cell * ELL width + slot) */
  bool ghosts_from_caller = false; /* host-Vec apply: u_local arrives with its ghosts filled (DMGlobalToLocal already ran) */
  /* tiled-kernel device arrays */
  TiledDev td;
  bool has_tiled = false;
  const char *last_kernel = "none";
  /* device-side time loop (rdyb_operator_advance_device): control block, its pinned mirror, the instantiated graph */
  struct {
    AdvCtl *ctl = nullptr, *mirror = nullptr;
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    cudaEvent_t ev_mirror = nullptr;
    bool dirty = false;          /* a launched interval has not been reflected in `cur` / the host's view of t, dt yet */
    /* what the graph was built for: rebuilt when any of it changes */
    double interval = 0.0, target = 0.0, maxinc = 0.0;
    int adaptive = -1, kernel_kind = -1, tile_cells = -1, opts = -1, max_ctas = -1;
    const double *src_sig = nullptr;
  } adv;
  /* Bradford-Sanders well balancing (libCEED semantics, rdyb_operator_enable_bs2002) */
  struct {
    bool enabled = false;
    int32_t nv = 0;
    double *vz = nullptr, *eta = nullptr, *cz = nullptr; /* vertex z, vertex eta, sorted vertex z per device cell [3][stride] */
    int32_t *vc_off = nullptr, *vc_cell = nullptr, *vb = nullptr, *ve = nullptr;
  } bs;
  const AdvCtl *adv_ctl_arg = nullptr; /* set around the launches captured into the device-side loop: fill_args passes it on */
  int adv_buf_arg = 0;
  int max_ctas = 0; /* cap on the grid of the persistent kernels (0 = none): rdyb_operator_set_max_ctas */
  long long p2p_timeout_cycles = 40000000000ll; /* in-kernel / wait-kernel neighbour time-out (~20 s; RDYB_P2P_TIMEOUT_S) */
  unsigned long long *wait_cycles = nullptr; /* [1] clock64() ticks the producers of this rank spent waiting for neighbours (sum over CTAs) */
  int kernel_kind = 0; /* 0 = gather (K1), 1 = tiled (K2), 2 = tiled + bulk-async staging (K3), 3 = K3 with compile-time tile shapes (K4; falls back to K3) */
  /* options */
  bool opt_courant = true, opt_fdiv = false, opt_prim = false, opt_bflux = false;
  bool mom_src = false; /* any momentum external source set */
  /* peer-memory halo (Euler fast path) */
  bool p2p = false;
  bool wait_in_kernel = false; /* set around the one-launch p2p step: K4 waits for the neighbours inside the kernel */
  PeerTable ptab;
  unsigned int p2p_epoch = 0;
  unsigned int *p2p_flags = nullptr; /* [RDYB_MAX_PEERS] counters raised by the neighbours, + 1 status word */
  void *p2p_opened[4 * RDYB_MAX_PEERS];
  int p2p_nopened = 0;
  /* fused push (K4 stores the border cells of each finished frontier tile into the neighbours itself; swe_tiled_fixed.cuh) */
  struct {
    bool built = false, active = false;
    int2 *range = nullptr;      /* [num_tiles] send entries of each tile */
    int32_t *src = nullptr, *rs = nullptr, *order = nullptr;
    double **dst[2] = {nullptr, nullptr}; /* per ping-pong buffer */
    unsigned int *counter = nullptr;
    unsigned int send_tiles = 0, launches = 0;
  } fpush;
  /* comm */
  ncclComm_t comm = nullptr;
  int32_t nranks = 1, rank = 0;
  /* timing of the last step_euler call */
  double last_ms = 0.0;
  int32_t last_launches = 0;
  /* pipelined host-Vec apply (rdyb_operator_apply_host): u_local is uploaded in chunks of view cells; a tile runs as soon as every
   * cell it reads has arrived, and a range of F goes back to the host as soon as every tile that writes into it has run, so the
   * host->device and device->host copies overlap (PCIe is full duplex) instead of running one after the other */
  struct HostPipe {
    bool built = false, usable = false;
    int32_t tile_cells = 0, nchunks = 0, chunk_cells = 0, out_cells = 0;
    int32_t *order = nullptr, *local_to_dev = nullptr, *owned_to_dev = nullptr; /* device */
    double *stage_in = nullptr, *stage_out = nullptr;                           /* device */
    std::vector<int32_t> gi, gf, out_need; /* launch-order group offsets (interior / frontier tiles), group an F range waits for */
    cudaStream_t s_in = nullptr, s_out = nullptr;
    std::vector<cudaEvent_t> ev_in, ev_out;
  } pipe;
};

template <typename T>
static int dalloc(T **p, size_t n) {
  CU(cudaMalloc((void **)p, sizeof(T) * (n > 0 ? n : 1)));
  CU(cudaMemset(*p, 0, sizeof(T) * (n > 0 ? n : 1)));
  return RDYB_SUCCESS;
}
template <typename T>
static int upload(T **p, const T *src, size_t n) {
  RDYB_CALL(dalloc(p, n));
  if (n) CU(cudaMemcpy(*p, src, sizeof(T) * n, cudaMemcpyHostToDevice));
  return RDYB_SUCCESS;
}

extern "C" int64_t rdyb_kernel_launch_count(void) { return g_launches.load(); }
extern "C" int64_t rdyb_host_sync_count(void) { return g_host_syncs.load(); }

static int grid_for(int64_t n, int block) { return (int)((n + block - 1) / block); }

static void adv_release_graph(RdybOperator op);
static int adv_resolve(RdybOperator op);
static void fpush_release(RdybOperator op) { /* the send lists depend on the tiling and on the peer table */
  void *fp[] = {op->fpush.range, op->fpush.src, op->fpush.rs, op->fpush.order, op->fpush.dst[0], op->fpush.dst[1], op->fpush.counter};
  for (void *q : fp)
    if (q) cudaFree(q);
  op->fpush = {};
}

extern "C" int rdyb_operator_destroy(RdybOperator *pop) {
  if (!pop || !*pop) return RDYB_SUCCESS;
  RdybOperator op = *pop;
  cudaSetDevice(op->device);
  if (op->stream) rdyb_stream_sync(op->stream);
  if (op->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(op->comm);
  for (int i = 0; i < op->p2p_nopened; ++i) cudaIpcCloseMemHandle(op->p2p_opened[i]);
  for (void *q : {(void *)op->bs.vz, (void *)op->bs.eta, (void *)op->bs.cz, (void *)op->bs.vc_off, (void *)op->bs.vc_cell, (void *)op->bs.vb, (void *)op->bs.ve})
    if (q) cudaFree(q);
  adv_release_graph(op);
  if (op->adv.ctl) cudaFree(op->adv.ctl);
  if (op->adv.mirror) cudaFreeHost(op->adv.mirror);
  if (op->adv.ev_mirror) cudaEventDestroy(op->adv.ev_mirror);
  if (op->p2p_flags) cudaFree(op->p2p_flags);
  if (op->wait_cycles) cudaFree(op->wait_cycles);
  fpush_release(op);
  void *ptrs[] = {op->state[0], op->state[1], op->F, op->src, op->mann, op->fdiv, op->prim, op->inv_area, op->area, op->dzdx, op->dzdy, op->zc,
                  op->ell_nbr, op->ell_sn, op->ell_cn, op->ell_slen, op->btype, op->bval, op->bflux, op->bacc, op->dev_to_local,
                  op->dev_to_owned, op->send_cells, op->stage, op->stage_idx, op->sendbuf, op->recvbuf, op->cmax, op->carg, op->partial,
                  op->rk_k[0], op->rk_k[1], op->rk_k[2], op->rk_k[3], op->rk_y, op->grad, op->m_gcx, op->m_gcy, op->m_dxs, op->m_dys,
                  op->m_dxn, op->m_dyn, op->t_edge_off, op->t_halo2, op->t_halo_nbr, op->t_halo_rec};
  for (void *p : ptrs)
    if (p) cudaFree(p);
  for (void *p : {(void *)op->pipe.order, (void *)op->pipe.local_to_dev, (void *)op->pipe.owned_to_dev, (void *)op->pipe.stage_in, (void *)op->pipe.stage_out})
    if (p) cudaFree(p);
  for (cudaEvent_t e : op->pipe.ev_in) cudaEventDestroy(e);
  for (cudaEvent_t e : op->pipe.ev_out) cudaEventDestroy(e);
  if (op->pipe.s_in) cudaStreamDestroy(op->pipe.s_in);
  if (op->pipe.s_out) cudaStreamDestroy(op->pipe.s_out);
  tiled_dev_free(&op->td);
  if (op->ev_pack) cudaEventDestroy(op->ev_pack);
  if (op->ev_halo) cudaEventDestroy(op->ev_halo);
  if (op->ev_t0) cudaEventDestroy(op->ev_t0);
  if (op->ev_t1) cudaEventDestroy(op->ev_t1);
  if (op->comm_stream) cudaStreamDestroy(op->comm_stream);
  if (op->own_stream && op->stream) cudaStreamDestroy(op->stream);
  rdyb_layout_free(&op->L);
  rdyb_tiled_layout_free(&op->T);
  delete op;
  *pop = nullptr;
  return RDYB_SUCCESS;
}

static int operator_create_impl(RdybOperator op, const RdybConfig *config, const RdybMeshView *mesh, int32_t nb, const RdybBoundaryView *bnd,
                                const RdybHaloView *halo, int device, void *cuda_stream) {
  op->cfg = *config;
  op->device = device;
  RDYB_CHECK(config->source_method == RDYB_SOURCE_SEMI_IMPLICIT || config->source_method == RDYB_SOURCE_IMPLICIT_XQ2018, RDYB_ERR_USER,
             "Only semi_implicit and implicit_xq2018 are supported in the PETSc version"); /* src/swe/swe_petsc.c:973-975 */
  RDYB_CHECK(config->well_balancing != RDYB_WELL_BALANCING_HR || mesh->cell_zc, RDYB_ERR_USER, "hydrostatic reconstruction needs cell_zc");
  RDYB_CHECK(config->reconstruction >= RDYB_RECON_FIRST_ORDER && config->reconstruction <= RDYB_RECON_MUSCL_VANLEER, RDYB_ERR_USER,
             "Invalid reconstruction: %d", config->reconstruction);
  /* src/operator.c:388-389 */
  RDYB_CHECK(!config->reconstruction || config->well_balancing != RDYB_WELL_BALANCING_HR, RDYB_ERR_USER,
             "-second_order cannot be used with well_balancing = HR simultaneously (not yet implemented)");
  op->recon = config->reconstruction;
  RDYB_CHECK(!op->recon || (mesh->cell_xc && mesh->cell_yc && mesh->edge_xm && mesh->edge_ym), RDYB_ERR_USER,
             "second-order reconstruction needs cell centroids and edge midpoints in the mesh view");
  for (int32_t b = 0; b < nb; ++b) {
    const int t = bnd[b].bc_type;
    RDYB_CHECK(t == RDYB_CONDITION_DIRICHLET || t == RDYB_CONDITION_REFLECTING || t == RDYB_CONDITION_CRITICAL_OUTFLOW, RDYB_ERR_USER,
               "Invalid boundary condition encountered for boundary %d\n", bnd[b].id); /* src/swe/swe_petsc.c:567-569 */
  }
  int ndev = 0;
  cudaError_t ce = cudaGetDeviceCount(&ndev);
  if (ce != cudaSuccess || ndev == 0)
    return rdyb_fail(RDYB_ERR_CUDA, "no CUDA device available (%s): the SWE path has no CPU fallback", cudaGetErrorString(ce));
  RDYB_CHECK(device >= 0 && device < ndev, RDYB_ERR_ARG, "device %d out of range (%d devices)", device, ndev);
  CU(cudaSetDevice(device));
  const char *no_reorder = getenv("RDYB_NO_REORDER");
  RDYB_CALL(rdyb_layout_build(mesh, nb, bnd, halo, !(no_reorder && no_reorder[0] == '1'), &op->L));
  const RdybLayout &L = op->L;
  op->N = L.n_owned; op->NL = L.n_local; op->NI = L.n_interior; op->stride = L.stride; op->width = L.width; op->nbe = L.num_boundary_edges;
  if (cuda_stream) op->stream = (cudaStream_t)cuda_stream;
  else { CU(cudaStreamCreateWithFlags(&op->stream, cudaStreamNonBlocking)); op->own_stream = true; }
  CU(cudaStreamCreateWithFlags(&op->comm_stream, cudaStreamNonBlocking));
  CU(cudaEventCreateWithFlags(&op->ev_pack, cudaEventDisableTiming));
  CU(cudaEventCreateWithFlags(&op->ev_halo, cudaEventDisableTiming));
  CU(cudaEventCreate(&op->ev_t0));
  CU(cudaEventCreate(&op->ev_t1));
  const size_t S = (size_t)L.stride;
  RDYB_CALL(dalloc(&op->state[0], 3 * S)); RDYB_CALL(dalloc(&op->state[1], 3 * S)); RDYB_CALL(dalloc(&op->F, 3 * S));
  RDYB_CALL(dalloc(&op->src, 3 * S)); RDYB_CALL(dalloc(&op->mann, S));
  /* the staging buffer also receives u_local in the VIEW's cell order, which may hold cells the layout dropped (non-face ghosts) */
  RDYB_CALL(dalloc(&op->stage, 3 * (S > (size_t)L.num_view_cells ? S : (size_t)L.num_view_cells)));
  RDYB_CALL(dalloc(&op->stage_idx, S));
  RDYB_CALL(upload(&op->inv_area, L.inv_area, S)); RDYB_CALL(upload(&op->area, L.area, S));
  RDYB_CALL(upload(&op->dzdx, L.dzdx, S)); RDYB_CALL(upload(&op->dzdy, L.dzdy, S));
  if (L.zc) RDYB_CALL(upload(&op->zc, L.zc, S));
  const size_t W = (size_t)L.width * S;
  RDYB_CALL(upload(&op->ell_nbr, L.ell_nbr, W)); RDYB_CALL(upload(&op->ell_sn, L.ell_sn, W));
  RDYB_CALL(upload(&op->ell_cn, L.ell_cn, W)); RDYB_CALL(upload(&op->ell_slen, L.ell_slen, W));
  RDYB_CALL(upload(&op->btype, L.b_type, (size_t)L.num_boundary_edges));
  RDYB_CALL(dalloc(&op->bval, 3 * (size_t)L.num_boundary_edges)); RDYB_CALL(dalloc(&op->bflux, 3 * (size_t)L.num_boundary_edges));
  RDYB_CALL(dalloc(&op->bacc, 3 * (size_t)L.num_boundary_edges));
  RDYB_CALL(upload(&op->dev_to_local, L.dev_to_local, (size_t)L.n_local)); RDYB_CALL(upload(&op->dev_to_owned, L.dev_to_owned, (size_t)L.n_owned));
  if (L.num_peers > 0) {
    const size_t ns = (size_t)L.send_offsets[L.num_peers], nr = (size_t)L.recv_offsets[L.num_peers];
    RDYB_CALL(upload(&op->send_cells, L.send_cells, ns));
    RDYB_CALL(dalloc(&op->sendbuf, 3 * ns)); RDYB_CALL(dalloc(&op->recvbuf, 3 * nr));
  }
  op->ncblocks_interior = grid_for(op->NI, RDYB_BLOCK);
  op->ncblocks = op->ncblocks_interior + grid_for(op->N - op->NI, RDYB_BLOCK) + 1;
  /* the tiled kernels need one slot per tile (K2, K2M) or per resident block: set_kernel grows the arrays when a tiled layout has
   * more tiles than this (short frontier tiles) */
  RDYB_CALL(dalloc(&op->cmax, (size_t)op->ncblocks)); RDYB_CALL(dalloc(&op->carg, (size_t)op->ncblocks));
  CU(cudaMemset(op->carg, 0xff, sizeof(long long) * (size_t)op->ncblocks));
  RDYB_CALL(dalloc(&op->partial, 1024));
  if (op->recon) {
    RDYB_CALL(rdyb_layout_build_muscl(mesh, &op->L));
    RDYB_CALL(dalloc(&op->grad, 6 * S));
    RDYB_CALL(upload(&op->m_gcx, L.ell_gcx, W)); RDYB_CALL(upload(&op->m_gcy, L.ell_gcy, W));
    RDYB_CALL(upload(&op->m_dxs, L.ell_dxs, W)); RDYB_CALL(upload(&op->m_dys, L.ell_dys, W));
    RDYB_CALL(upload(&op->m_dxn, L.ell_dxn, W)); RDYB_CALL(upload(&op->m_dyn, L.ell_dyn, W));
  }
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_create(const RdybConfig *config, const RdybMeshView *mesh, int32_t nb, const RdybBoundaryView *bnd,
                                    const RdybHaloView *halo, int device, void *cuda_stream, RdybOperator *out) {
  RDYB_CHECK(config && mesh && out, RDYB_ERR_ARG, "rdyb_operator_create: NULL argument");
  RDYB_CHECK(nb >= 0 && (nb == 0 || bnd), RDYB_ERR_ARG, "rdyb_operator_create: %d boundaries but no boundary array", nb);
  RdybOperator op = new RdybOperator_();
  memset(&op->L, 0, sizeof(op->L));
  memset(&op->T, 0, sizeof(op->T));
  memset(&op->td, 0, sizeof(op->td));
  memset(&op->ptab, 0, sizeof(op->ptab));
  int ierr = operator_create_impl(op, config, mesh, nb, bnd, halo, device, cuda_stream);
  if (ierr) {
    char saved[1024];
    snprintf(saved, sizeof(saved), "%s", rdyb_last_error());
    rdyb_operator_destroy(&op);
    return rdyb_fail(ierr, "%s", saved);
  }
  *out = op;
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_num_owned_cells(RdybOperator op, int32_t *n) {
  RDYB_CHECK(op && n, RDYB_ERR_ARG, "NULL argument");
  *n = op->N;
  return RDYB_SUCCESS;
}
extern "C" int rdyb_operator_num_local_cells(RdybOperator op, int32_t *n) {
  RDYB_CHECK(op && n, RDYB_ERR_ARG, "NULL argument");
  *n = op->NL;
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_set_options(RdybOperator op, int courant, int flux_divergence, int primitive_variables, int boundary_fluxes) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "NULL operator");
  CU(cudaSetDevice(op->device));
  op->opt_courant = courant != 0; op->opt_fdiv = flux_divergence != 0; op->opt_prim = primitive_variables != 0; op->opt_bflux = boundary_fluxes != 0;
  if (op->opt_fdiv && !op->fdiv) RDYB_CALL(dalloc(&op->fdiv, 3 * (size_t)op->stride));
  if (op->opt_prim && !op->prim) RDYB_CALL(dalloc(&op->prim, 3 * (size_t)op->stride));
  return RDYB_SUCCESS;
}

/* at most n CTAs in the grid of the persistent kernels (0 = no cap).  On a small mesh this makes every CTA walk many tiles, i.e.
 * all pipeline stages and both mbarrier parities: the regime of the 10 M-cell bench, reachable by the parity tests. */
extern "C" int rdyb_operator_set_max_ctas(RdybOperator op, int32_t n) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "NULL operator");
  RDYB_CHECK(n >= 0, RDYB_ERR_ARG, "negative CTA cap");
  op->max_ctas = n;
  return RDYB_SUCCESS;
}

static void pipe_release(RdybOperator op);
/* one Courant slot per tile for the kernels that index them by tile number */
static int ensure_courant_slots(RdybOperator op, int32_t n) {
  if (n <= op->ncblocks) return RDYB_SUCCESS;
  CU(rdyb_stream_sync(op->stream));
  if (op->cmax) cudaFree(op->cmax);
  if (op->carg) cudaFree(op->carg);
  op->cmax = nullptr; op->carg = nullptr;
  op->ncblocks = n;
  RDYB_CALL(dalloc(&op->cmax, (size_t)n)); RDYB_CALL(dalloc(&op->carg, (size_t)n));
  CU(cudaMemset(op->carg, 0xff, sizeof(long long) * (size_t)n));
  return RDYB_SUCCESS;
}
/* kernel selection: 0 = K1 gather, 1 = K2 tiled, 2 = K3 bulk-async staging, 3 = K4 compile-time tile shapes (the tiled layout is built on first use) */
extern "C" int rdyb_operator_set_kernel(RdybOperator op, int32_t kind, int32_t tile_cells) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "NULL operator");
  RDYB_CHECK(kind >= 0 && kind <= 3, RDYB_ERR_ARG, "unknown kernel kind %d", kind);
  CU(cudaSetDevice(op->device));
  if (kind >= 1) {
    if (op->has_tiled && op->T.tile_cells != tile_cells) {
      tiled_dev_free(&op->td);
      rdyb_tiled_layout_free(&op->T);
      if (op->t_edge_off) { cudaFree(op->t_edge_off); op->t_edge_off = nullptr; }
      if (op->t_halo2) { cudaFree(op->t_halo2); op->t_halo2 = nullptr; }
      if (op->t_halo_nbr) { cudaFree(op->t_halo_nbr); op->t_halo_nbr = nullptr; }
      if (op->t_halo_rec) { cudaFree(op->t_halo_rec); op->t_halo_rec = nullptr; }
      op->has_tiled = false;
      pipe_release(op);
      fpush_release(op);
    }
    if (!op->has_tiled) {
      RDYB_CALL(rdyb_tiled_layout_build(&op->L, tile_cells, &op->T));
      const bool muscl_tiles = op->recon && !getenv("RDYB_MUSCL_K1"); /* second order on the tiled layout; RDYB_MUSCL_K1 keeps it on K1 */
      if (muscl_tiles) RDYB_CALL(rdyb_tiled_layout_build_muscl(&op->L, &op->T)); /* BEFORE the upload: it completes the tile descriptors (second ring) */
      RDYB_CALL(tiled_dev_upload(&op->T, &op->L, &op->td));
      RDYB_CALL(ensure_courant_slots(op, op->T.num_tiles + 1));
      if (muscl_tiles) {
        RDYB_CALL(upload(&op->t_edge_off, op->T.edge_off, 4 * (size_t)op->T.edge_off_plane));
        if (op->T.two_ring) {
          RDYB_CALL(upload(&op->t_halo2, op->T.halo2_cells, (size_t)op->T.total_halo2));
          RDYB_CALL(upload(&op->t_halo_nbr, op->T.halo_nbr, 4 * (size_t)op->T.total_halo));
          if (op->L.width == 3) {
            std::vector<double> rec(8 * (size_t)op->N, 0.0);
            const size_t S = (size_t)op->L.stride;
            for (int32_t c = 0; c < op->N; ++c) {
              double *r = &rec[8 * (size_t)c];
              r[0] = op->L.inv_area[c];
              for (int k = 0; k < 3; ++k) { r[1 + k] = op->L.ell_gcx[(size_t)k * S + c]; r[4 + k] = op->L.ell_gcy[(size_t)k * S + c]; }
            }
            RDYB_CALL(upload(&op->t_halo_rec, rec.data(), rec.size()));
          }
        }
      }
      op->has_tiled = true;
    }
  }
  op->kernel_kind = kind;
  return RDYB_SUCCESS;
}


/* peer-memory halo: a neighbour that did not deliver its border cells before the time-out leaves a status word behind; the step
 * then ran on stale ghosts, so every call that hands data back to the host fails instead of returning it.  Call after a sync. */
static int p2p_check_status(RdybOperator op) {
  if (!op->p2p || !op->p2p_flags) return RDYB_SUCCESS;
  int status = 0;
  CU(cudaMemcpy(&status, op->p2p_flags + RDYB_MAX_PEERS, sizeof(int), cudaMemcpyDeviceToHost));
  if (status) {
    const int pidx = status - 1;
    return rdyb_fail(RDYB_ERR_NCCL, "peer-memory halo: neighbour %d (rank %d) did not deliver its border cells in time; the state of this rank is invalid",
                     pidx, pidx >= 0 && pidx < op->L.num_peers ? op->L.peer_ranks[pidx] : -1);
  }
  return RDYB_SUCCESS;
}

/* ---------------- data setters / getters ---------------- */
extern "C" int rdyb_operator_set_boundary_values(RdybOperator op, int32_t b, int32_t comp_offset, int32_t num_comp, int32_t num_edges,
                                                 const double *values) {
  RDYB_CHECK(op && values, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(b >= 0 && b < op->L.num_boundaries, RDYB_ERR_USER, "Invalid boundary index: %d", b);
  const int32_t ne = op->L.b_offsets[b + 1] - op->L.b_offsets[b];
  /* src/operator.c:1050-1052 */
  RDYB_CHECK(num_edges == ne, RDYB_ERR_USER, "Wrong number of boundary edges for boundary %d: %d (should be %d)", b, num_edges, ne);
  RDYB_CHECK(comp_offset >= 0 && num_comp > 0 && comp_offset + num_comp <= 3, RDYB_ERR_USER, "Invalid component range [%d, %d)", comp_offset,
             comp_offset + num_comp);
  CU(cudaSetDevice(op->device));
  std::vector<double> tmp((size_t)ne);
  for (int32_t c = 0; c < num_comp; ++c) {
    for (int32_t e = 0; e < ne; ++e) tmp[e] = values[(size_t)num_comp * e + c];
    CU(cudaMemcpyAsync(op->bval + (size_t)(comp_offset + c) * op->nbe + op->L.b_offsets[b], tmp.data(), sizeof(double) * ne, cudaMemcpyHostToDevice,
                       op->stream));
    CU(rdyb_stream_sync(op->stream));
  }
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_get_boundary_fluxes(RdybOperator op, int32_t b, int32_t num_edges, int accumulated, double *values) {
  RDYB_CHECK(op && values, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(b >= 0 && b < op->L.num_boundaries, RDYB_ERR_USER, "Invalid boundary index: %d", b);
  const int32_t ne = op->L.b_offsets[b + 1] - op->L.b_offsets[b];
  RDYB_CHECK(num_edges == ne, RDYB_ERR_USER, "Wrong number of boundary edges for boundary %d: %d (should be %d)", b, num_edges, ne);
  RDYB_CHECK(op->opt_bflux, RDYB_ERR_USER, "boundary fluxes are not being recorded (rdyb_operator_set_options)");
  CU(cudaSetDevice(op->device));
  const double *srcp = accumulated ? op->bacc : op->bflux;
  for (int32_t c = 0; c < 3; ++c)
    CU(cudaMemcpyAsync(values + (size_t)c * ne, srcp + (size_t)c * op->nbe + op->L.b_offsets[b], sizeof(double) * ne, cudaMemcpyDeviceToHost, op->stream));
  CU(rdyb_stream_sync(op->stream));
  return RDYB_SUCCESS;
}

/* owned-order host array -> one device-order component array */
static int upload_owned_component(RdybOperator op, const double *values, double *dst) {
  CU(cudaMemcpyAsync(op->stage, values, sizeof(double) * op->N, cudaMemcpyHostToDevice, op->stream));
  gather_kernel<<<grid_for(op->N, 256), 256, 0, op->stream>>>(op->stage, op->dev_to_owned, dst, op->N);
  LAUNCHED();
  CU(cudaGetLastError());
  CU(rdyb_stream_sync(op->stream));
  return RDYB_SUCCESS;
}
static int download_owned_component(RdybOperator op, const double *srcp, double *values) {
  scatter_kernel<<<grid_for(op->N, 256), 256, 0, op->stream>>>(srcp, op->dev_to_owned, op->stage, op->N);
  LAUNCHED();
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(values, op->stage, sizeof(double) * op->N, cudaMemcpyDeviceToHost, op->stream));
  CU(rdyb_stream_sync(op->stream));
  return p2p_check_status(op);
}
/* values[i] -> dst[owned_to_dev[ids[i]]] */
static int upload_regional(RdybOperator op, int32_t n, const int32_t *ids, const double *values, double *dst) {
  std::vector<int32_t> idx((size_t)n);
  for (int32_t i = 0; i < n; ++i) {
    RDYB_CHECK(ids[i] >= 0 && ids[i] < op->N, RDYB_ERR_USER, "owned cell id %d out of range", ids[i]);
    idx[i] = op->L.owned_to_dev[ids[i]];
  }
  RDYB_CHECK(n <= op->stride, RDYB_ERR_ARG_SIZ, "region has more cells (%d) than the mesh", n);
  CU(cudaMemcpyAsync(op->stage, values, sizeof(double) * n, cudaMemcpyHostToDevice, op->stream));
  CU(cudaMemcpyAsync(op->stage_idx, idx.data(), sizeof(int32_t) * n, cudaMemcpyHostToDevice, op->stream));
  scatter_kernel<<<grid_for(n, 256), 256, 0, op->stream>>>(op->stage, op->stage_idx, dst, n);
  LAUNCHED();
  CU(cudaGetLastError());
  CU(rdyb_stream_sync(op->stream));
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_set_external_source(RdybOperator op, int32_t component, int32_t num_owned, const double *values) {
  RDYB_CHECK(op && values, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(component >= 0 && component < 3, RDYB_ERR_USER, "Invalid source component: %d", component);
  RDYB_CHECK(num_owned == op->N, RDYB_ERR_USER, "Wrong size (%d) for owned cell data (should be %d)", num_owned, op->N);
  CU(cudaSetDevice(op->device));
  if (component > 0) op->mom_src = true;
  return upload_owned_component(op, values, op->src + (size_t)component * op->stride);
}
/* the Operator's external_sources Vec as it is: AoS [num_owned][3] in owned order (src/operator.c:93-96) */
extern "C" int rdyb_operator_set_external_source_aos(RdybOperator op, int32_t num_owned, const double *values) {
  RDYB_CHECK(op && values, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(num_owned == op->N, RDYB_ERR_USER, "Wrong size (%d) for owned cell data (should be %d)", num_owned, op->N);
  CU(cudaSetDevice(op->device));
  bool mom = false;
  for (int64_t i = 0; i < op->N && !mom; ++i) mom = values[3 * i + 1] != 0.0 || values[3 * i + 2] != 0.0;
  op->mom_src = mom;
  CU(cudaMemcpyAsync(op->stage, values, sizeof(double) * 3 * op->N, cudaMemcpyHostToDevice, op->stream));
  aos_to_soa_kernel<<<grid_for(op->N, 256), 256, 0, op->stream>>>(op->stage, op->dev_to_owned, op->src, op->N, op->stride);
  LAUNCHED();
  CU(cudaGetLastError());
  CU(rdyb_stream_sync(op->stream));
  return RDYB_SUCCESS;
}
extern "C" int rdyb_operator_get_external_source(RdybOperator op, int32_t component, int32_t num_owned, double *values) {
  RDYB_CHECK(op && values, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(component >= 0 && component < 3, RDYB_ERR_USER, "Invalid source component: %d", component);
  RDYB_CHECK(num_owned == op->N, RDYB_ERR_USER, "Wrong size (%d) for owned cell data (should be %d)", num_owned, op->N);
  CU(cudaSetDevice(op->device));
  return download_owned_component(op, op->src + (size_t)component * op->stride, values);
}
extern "C" int rdyb_operator_set_regional_external_source(RdybOperator op, int32_t component, int32_t n, const int32_t *ids, const double *values) {
  RDYB_CHECK(op && (n == 0 || (ids && values)), RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(component >= 0 && component < 3, RDYB_ERR_USER, "Invalid source component: %d", component);
  CU(cudaSetDevice(op->device));
  if (component > 0) op->mom_src = true;
  return n ? upload_regional(op, n, ids, values, op->src + (size_t)component * op->stride) : RDYB_SUCCESS;
}
extern "C" int rdyb_operator_set_mannings(RdybOperator op, int32_t num_owned, const double *values) {
  RDYB_CHECK(op && values, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(num_owned == op->N, RDYB_ERR_USER, "Wrong size (%d) for owned cell data (should be %d)", num_owned, op->N);
  CU(cudaSetDevice(op->device));
  return upload_owned_component(op, values, op->mann);
}
extern "C" int rdyb_operator_set_regional_mannings(RdybOperator op, int32_t n, const int32_t *ids, const double *values) {
  RDYB_CHECK(op && (n == 0 || (ids && values)), RDYB_ERR_ARG, "NULL argument");
  CU(cudaSetDevice(op->device));
  return n ? upload_regional(op, n, ids, values, op->mann) : RDYB_SUCCESS;
}

extern "C" int rdyb_operator_set_state(RdybOperator op, const double *u_global) {
  RDYB_CHECK(op && u_global, RDYB_ERR_ARG, "NULL argument");
  CU(cudaSetDevice(op->device));
  RDYB_CALL(adv_resolve(op));
  CU(cudaMemcpyAsync(op->stage, u_global, sizeof(double) * 3 * op->N, cudaMemcpyHostToDevice, op->stream));
  aos_to_soa_kernel<<<grid_for(op->N, 256), 256, 0, op->stream>>>(op->stage, op->dev_to_owned, op->state[op->cur], op->N, op->stride);
  LAUNCHED();
  CU(cudaGetLastError());
  CU(rdyb_stream_sync(op->stream));
  return RDYB_SUCCESS;
}
extern "C" int rdyb_operator_get_state(RdybOperator op, double *u_global) {
  RDYB_CHECK(op && u_global, RDYB_ERR_ARG, "NULL argument");
  CU(cudaSetDevice(op->device));
  RDYB_CALL(adv_resolve(op));
  soa_to_aos_kernel<<<grid_for(op->N, 256), 256, 0, op->stream>>>(op->state[op->cur], op->dev_to_owned, op->stage, op->N, op->stride);
  LAUNCHED();
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(u_global, op->stage, sizeof(double) * 3 * op->N, cudaMemcpyDeviceToHost, op->stream));
  CU(rdyb_stream_sync(op->stream));
  return p2p_check_status(op);
}
extern "C" int rdyb_operator_get_state_component(RdybOperator op, int32_t component, int32_t num_owned, double *values) {
  RDYB_CHECK(op && values, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(component >= 0 && component < 3, RDYB_ERR_USER, "Invalid component: %d", component);
  RDYB_CHECK(num_owned == op->N, RDYB_ERR_USER, "Wrong size (%d) for owned cell data (should be %d)", num_owned, op->N);
  CU(cudaSetDevice(op->device));
  RDYB_CALL(adv_resolve(op));
  return download_owned_component(op, op->state[op->cur] + (size_t)component * op->stride, values);
}
/* optional output vectors in the reference's AoS owned layout: which = 0 flux_divergence, 1 primitive_variables, 2 F of the last apply */
extern "C" int rdyb_operator_get_field(RdybOperator op, int32_t which, double *values) {
  RDYB_CHECK(op && values, RDYB_ERR_ARG, "NULL argument");
  const double *srcp = which == 0 ? op->fdiv : which == 1 ? op->prim : which == 2 ? op->F : nullptr;
  RDYB_CHECK(srcp, RDYB_ERR_USER, "field %d is not being computed (rdyb_operator_set_options)", which);
  CU(cudaSetDevice(op->device));
  soa_to_aos_kernel<<<grid_for(op->N, 256), 256, 0, op->stream>>>(srcp, op->dev_to_owned, op->stage, op->N, op->stride);
  LAUNCHED();
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(values, op->stage, sizeof(double) * 3 * op->N, cudaMemcpyDeviceToHost, op->stream));
  CU(rdyb_stream_sync(op->stream));
  return p2p_check_status(op);
}

/* ---------------- launches ---------------- */
static void fill_args(RdybOperator op, SweArgs &a, const double *in, double *out, double dt) {
  const size_t S = (size_t)op->stride;
  a.h = in; a.hu = in + S; a.hv = in + 2 * S;
  a.o0 = out; a.o1 = out + S; a.o2 = out + 2 * S;
  a.inv_area = op->inv_area; a.dzdx = op->dzdx; a.dzdy = op->dzdy; a.zc = op->zc; a.mann = op->mann;
  a.s0 = op->src; a.s1 = op->mom_src ? op->src + S : nullptr; a.s2 = op->mom_src ? op->src + 2 * S : nullptr;
  a.nbr = op->ell_nbr; a.esn = op->ell_sn; a.ecn = op->ell_cn; a.eslen = op->ell_slen;
  a.btype = op->btype; a.bval = op->bval;
  a.bflux = op->opt_bflux ? op->bflux : nullptr; a.bacc = op->opt_bflux ? op->bacc : nullptr;
  a.nbe = op->nbe;
  a.fdiv = op->opt_fdiv ? op->fdiv : nullptr; a.prim = op->opt_prim ? op->prim : nullptr;
  a.cmax = nullptr; a.carg = nullptr;
  a.stride = op->stride; a.width = op->width; a.begin = 0; a.end = 0;
  a.grad = op->grad; a.gcx = op->m_gcx; a.gcy = op->m_gcy; a.dxs = op->m_dxs; a.dys = op->m_dys; a.dxn = op->m_dxn; a.dyn = op->m_dyn;
  a.ctl = op->adv_ctl_arg; a.ctl_buf = op->adv_buf_arg;
  if (op->bs.enabled) { a.bs_eta = op->bs.eta; a.bs_vz = op->bs.vz; a.bs_vb = op->bs.vb; a.bs_ve = op->bs.ve; }
  a.dt = dt; a.tiny_h = op->cfg.tiny_h; a.ha2 = op->cfg.h_anuga_regular * op->cfg.h_anuga_regular; a.xq_thr = op->cfg.xq2018_threshold;
}

template <bool HR, int SRC, bool EULER>
static void launch_gather_c(RdybOperator op, const SweArgs &a, int grid) {
  if (op->opt_courant) swe_gather_kernel<HR, SRC, EULER, true><<<grid, RDYB_BLOCK, 0, op->stream>>>(a);
  else swe_gather_kernel<HR, SRC, EULER, false><<<grid, RDYB_BLOCK, 0, op->stream>>>(a);
}
template <int SRC, bool EULER, int RECON>
static void launch_gather_r(RdybOperator op, const SweArgs &a, int grid) {
  if (op->opt_courant) swe_gather_kernel<false, SRC, EULER, true, RECON><<<grid, RDYB_BLOCK, 0, op->stream>>>(a);
  else swe_gather_kernel<false, SRC, EULER, false, RECON><<<grid, RDYB_BLOCK, 0, op->stream>>>(a);
}
template <bool EULER>
static void launch_gather(RdybOperator op, const SweArgs &a, int grid) {
  const bool hr = op->cfg.well_balancing == RDYB_WELL_BALANCING_HR;
  const bool xq = op->cfg.source_method == RDYB_SOURCE_IMPLICIT_XQ2018;
  if (op->recon) { /* second order (never HR) */
    switch (op->recon) {
      case RDYB_RECON_MUSCL_MINMOD: if (xq) launch_gather_r<1, EULER, RDYB_RECON_MUSCL_MINMOD>(op, a, grid); else launch_gather_r<0, EULER, RDYB_RECON_MUSCL_MINMOD>(op, a, grid); break;
      case RDYB_RECON_MUSCL_NONE: if (xq) launch_gather_r<1, EULER, RDYB_RECON_MUSCL_NONE>(op, a, grid); else launch_gather_r<0, EULER, RDYB_RECON_MUSCL_NONE>(op, a, grid); break;
      default: if (xq) launch_gather_r<1, EULER, RDYB_RECON_MUSCL_VANLEER>(op, a, grid); else launch_gather_r<0, EULER, RDYB_RECON_MUSCL_VANLEER>(op, a, grid); break;
    }
    return;
  }
  if (hr) { if (xq) launch_gather_c<true, 1, EULER>(op, a, grid); else launch_gather_c<true, 0, EULER>(op, a, grid); }
  else { if (xq) launch_gather_c<false, 1, EULER>(op, a, grid); else launch_gather_c<false, 0, EULER>(op, a, grid); }
}

/* one RHS / Euler pass over owned device cells [begin,end) */
static int launch_cells(RdybOperator op, const double *in, double *out, double dt, bool euler, int32_t begin, int32_t end, int32_t cslot0) {
  if (end <= begin) return RDYB_SUCCESS;
  tl_rdyb_max_ctas = op->max_ctas;
  op->courant_arg_tiled = false;
  if (op->recon && op->kernel_kind >= 1 && op->has_tiled && begin == 0 && end == op->N && tiled_muscl_applicable(op->T, op->t_edge_off)) {
    /* second order on the tiled layout: every tiled kind runs K2M (one launch over all tiles; the state and gradient halos are complete) */
    SweArgs t;
    fill_args(op, t, in, out, dt);
    t.begin = begin; t.end = end;
    if (op->opt_courant) { t.cmax = op->cmax; t.carg = op->carg; } /* one slot per tile */
    op->courant_arg_tiled = true;
    const int k4m = getenv("RDYB_MUSCL_K2M") ? -1 : tiled_muscl_fixed_mode(op->T, op->t_edge_off, t, op->t_halo2 && op->t_halo_nbr);
    if (k4m >= 0) { /* K4M: compile-time tile shape; mode 2 computes the gradients inside the kernel (rhs_pass then skips the gradient pass) */
      op->last_kernel = k4m == 2 ? "swe_tiled_muscl_fixed_kernel (K4M, MUSCL, in-kernel gradients over a two-ring halo)"
                                 : "swe_tiled_muscl_fixed_kernel (K4M, MUSCL) + muscl_gradient_kernel";
      RDYB_CALL(tiled_muscl_fixed_launch(op->td, op->T, op->cfg, t, op->t_edge_off, op->t_halo2, op->t_halo_nbr, op->t_halo_rec, k4m, euler, op->stream));
    } else if (!getenv("RDYB_MUSCL_K2M") && tiled_muscl_async_applicable(op->T, op->t_edge_off)) { /* K3M: 15 % faster than K2M on the B200 (profiles/r2_bench_so_*.json); RDYB_MUSCL_K2M=1 keeps K2M */
      op->last_kernel = "swe_tiled_muscl_async_kernel (K3M, MUSCL) + muscl_gradient_kernel";
      RDYB_CALL(tiled_muscl_async_launch(op->td, op->T, op->cfg, t, op->t_edge_off, euler, op->stream));
    } else {
      op->last_kernel = "swe_tiled_muscl_kernel (K2M, MUSCL) + muscl_gradient_kernel";
      RDYB_CALL(tiled_muscl_launch(op->td, op->T, op->cfg, t, op->t_edge_off, euler, op->stream));
    }
    LAUNCHED();
    CU(cudaGetLastError());
    return RDYB_SUCCESS;
  }
  if (op->kernel_kind >= 1 && op->has_tiled && !op->recon && !op->bs.enabled) { /* first order: the tiled kernels (the BS2002 term lives in K1 only) */
    op->courant_arg_tiled = true;
    /* interior tiles and frontier tiles (thin strips along the partition border, with many more outside neighbours) are
     * launched separately so that each part can use the most specialised kernel its tiles fit */
    SweArgs t;
    fill_args(op, t, in, out, dt);
    /* frontier tiles are short (device_layout.c), so they normally fit K4's compiled shapes too: one launch for both parts */
    const bool whole_fits = op->kernel_kind == 3 && tiled_fixed_applicable(op->T.tile_cells, op->T.max_halo, op->T.max_edges, t);
    if (begin < op->NI && end > op->NI && !whole_fits) {
      RDYB_CALL(launch_cells(op, in, out, dt, euler, begin, op->NI, cslot0));
      return launch_cells(op, in, out, dt, euler, op->NI, end, cslot0);
    }
    t.begin = begin; t.end = end;
    if (op->opt_courant) { t.cmax = op->cmax; t.carg = op->carg; } /* one slot per tile, indexed by tile number */
    const bool interior = end <= op->NI;
    const int mh = interior ? op->T.max_halo_interior : op->T.max_halo, me = interior ? op->T.max_edges_interior : op->T.max_edges;
    FixedWait fw = {op->ptab.num_peers, op->p2p_epoch, op->p2p_flags, reinterpret_cast<int *>(op->p2p_flags + RDYB_MAX_PEERS)};
    fw.timeout = op->p2p_timeout_cycles; fw.wait_cycles = op->wait_cycles;
    if (op->wait_in_kernel && op->fpush.active && euler) { /* this launch also sends: the epoch the neighbours wait for next */
      const int buf = out == op->state[1];
      fw.order = op->fpush.order; fw.push_range = op->fpush.range; fw.push_src = op->fpush.src; fw.push_dst = op->fpush.dst[buf]; fw.push_rs = op->fpush.rs;
      fw.push_counter = op->fpush.counter; fw.push_target = ++op->fpush.launches * op->fpush.send_tiles;
      fw.push_epoch = op->p2p_epoch + 1; fw.push_flag = op->ptab.flag;
    }
    const FixedWait *wait = op->wait_in_kernel ? &fw : nullptr;
    if (op->kernel_kind >= 3 && tiled_fixed_applicable(op->T.tile_cells, mh, me, t)) {
      if (begin < op->NI || op->NI == 0) op->last_kernel = "swe_tiled_fixed_kernel (K4)";
      RDYB_CALL(tiled_fixed_launch(op->td, op->T, op->cfg, t, op->NI, euler, op->opt_courant, mh, me, op->stream, wait));
    } else if (op->kernel_kind >= 2) {
      if (interior || op->NI == 0) op->last_kernel = "swe_tiled_async_kernel (K3)";
      RDYB_CALL(tiled_async_launch(op->td, op->T, op->cfg, t, op->NI, euler, op->opt_courant, op->stream));
    } else {
      op->last_kernel = "swe_tiled_kernel (K2)";
      RDYB_CALL(tiled_launch(op->td, op->T, op->cfg, t, op->NI, euler, op->opt_courant, op->stream));
    }
    LAUNCHED();
    CU(cudaGetLastError());
    return RDYB_SUCCESS;
  }
  SweArgs a;
  fill_args(op, a, in, out, dt);
  a.begin = begin; a.end = end;
  if (op->opt_courant) { a.cmax = op->cmax + cslot0; a.carg = op->carg + cslot0; }
  op->last_kernel = op->recon ? "swe_gather_kernel (K1, MUSCL) + muscl_gradient_kernel" : "swe_gather_kernel (K1)";
  const int grid = grid_for(end - begin, RDYB_BLOCK);
  if (euler) launch_gather<true>(op, a, grid);
  else launch_gather<false>(op, a, grid);
  LAUNCHED();
  CU(cudaGetLastError());
  return RDYB_SUCCESS;
}

/* ---------------- halo exchange ---------------- */
static int halo_begin(RdybOperator op, double *state) {
  const RdybLayout &L = op->L;
  if (L.num_peers == 0 || !op->comm) return RDYB_SUCCESS;
  const int32_t ns = L.send_offsets[L.num_peers], nr = L.recv_offsets[L.num_peers];
  if (ns > 0) {
    halo_pack_kernel<<<grid_for(ns, 256), 256, 0, op->stream>>>(state, op->send_cells, op->sendbuf, ns, op->stride);
    LAUNCHED();
    CU(cudaGetLastError());
  }
  CU(cudaEventRecord(op->ev_pack, op->stream));
  CU(cudaStreamWaitEvent(op->comm_stream, op->ev_pack, 0));
  NC(g_nccl.GroupStart());
  for (int32_t p = 0; p < L.num_peers; ++p) {
    const int32_t s0 = L.send_offsets[p], sn = L.send_offsets[p + 1] - s0, r0 = L.recv_offsets[p], rn = L.recv_offsets[p + 1] - r0;
    for (int c = 0; c < 3; ++c) {
      /* one message per component so that both ends can address plain [3][n] buffers */
      if (sn > 0) NC(g_nccl.Send(op->sendbuf + (size_t)c * ns + s0, (size_t)sn, ncclDouble, L.peer_ranks[p], op->comm, op->comm_stream));
      if (rn > 0) NC(g_nccl.Recv(op->recvbuf + (size_t)c * nr + r0, (size_t)rn, ncclDouble, L.peer_ranks[p], op->comm, op->comm_stream));
    }
  }
  NC(g_nccl.GroupEnd());
  if (nr > 0) {
    halo_unpack_kernel<<<grid_for(nr, 256), 256, 0, op->comm_stream>>>(op->recvbuf, state, nr, op->N, op->stride);
    LAUNCHED();
    CU(cudaGetLastError());
  }
  CU(cudaEventRecord(op->ev_halo, op->comm_stream));
  return RDYB_SUCCESS;
}
static int halo_end(RdybOperator op) {
  if (op->L.num_peers == 0 || !op->comm) return RDYB_SUCCESS;
  CU(cudaStreamWaitEvent(op->stream, op->ev_halo, 0));
  return RDYB_SUCCESS;
}

extern "C" int rdyb_nccl_unique_id(void *id128) {
  RDYB_CHECK(id128, RDYB_ERR_ARG, "NULL argument");
  RDYB_CALL(nccl_load());
  ncclUniqueId id;
  NC(g_nccl.GetUniqueId(&id));
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is expected to be 128 bytes");
  memcpy(id128, &id, 128);
  return RDYB_SUCCESS;
}
extern "C" int rdyb_operator_init_comm(RdybOperator op, int32_t nranks, int32_t rank, const void *id128) {
  RDYB_CHECK(op && id128, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(nranks >= 1 && rank >= 0 && rank < nranks, RDYB_ERR_ARG, "bad rank %d of %d", rank, nranks);
  RDYB_CALL(nccl_load());
  CU(cudaSetDevice(op->device));
  ncclUniqueId id;
  memcpy(&id, id128, 128);
  NC(g_nccl.CommInitRank(&op->comm, nranks, id, rank));
  op->nranks = nranks; op->rank = rank;
  return RDYB_SUCCESS;
}
extern "C" int rdyb_operator_halo_exchange(RdybOperator op) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "NULL operator");
  CU(cudaSetDevice(op->device));
  RDYB_CALL(adv_resolve(op));
  RDYB_CALL(halo_begin(op, op->state[op->cur]));
  RDYB_CALL(halo_end(op));
  return RDYB_SUCCESS;
}


/* ---------------- peer-memory halo: export / import of the CUDA IPC handles ---------------- */
struct HaloBlob { /* <= RDYB_HALO_BLOB_BYTES */
  uint32_t magic;
  int32_t rank, n_owned, stride, num_peers;
  int32_t peer_ranks[RDYB_MAX_PEERS];
  int32_t recv_offsets[RDYB_MAX_PEERS + 1];
  int32_t send_counts[RDYB_MAX_PEERS];
  cudaIpcMemHandle_t state[3], flags; /* state[2]: the RK stage buffer (rk_y), so that RK4 stages can use the peer-memory halo too */
};
static_assert(sizeof(HaloBlob) <= RDYB_HALO_BLOB_BYTES, "halo blob does not fit its public size");

extern "C" int rdyb_operator_halo_export(RdybOperator op, int32_t rank, void *blob) {
  RDYB_CHECK(op && blob, RDYB_ERR_ARG, "NULL argument");
  const RdybLayout &L = op->L;
  RDYB_CHECK(L.num_peers <= RDYB_MAX_PEERS, RDYB_ERR_SUP, "peer-memory halo supports at most %d neighbours per rank (this rank has %d)", RDYB_MAX_PEERS, L.num_peers);
  CU(cudaSetDevice(op->device));
  if (!op->p2p_flags) RDYB_CALL(dalloc(&op->p2p_flags, RDYB_MAX_PEERS + 8));
  HaloBlob b;
  memset(&b, 0, sizeof(b));
  b.magic = 0x52445942u; b.rank = rank; b.n_owned = op->N; b.stride = op->stride; b.num_peers = L.num_peers;
  for (int p = 0; p < L.num_peers; ++p) { b.peer_ranks[p] = L.peer_ranks[p]; b.recv_offsets[p] = L.recv_offsets[p]; b.send_counts[p] = L.send_offsets[p + 1] - L.send_offsets[p]; }
  if (L.num_peers > 0) b.recv_offsets[L.num_peers] = L.recv_offsets[L.num_peers];
  if (!op->rk_y) RDYB_CALL(dalloc(&op->rk_y, 3 * (size_t)op->stride));
  CU(cudaIpcGetMemHandle(&b.state[0], op->state[0]));
  CU(cudaIpcGetMemHandle(&b.state[1], op->state[1]));
  CU(cudaIpcGetMemHandle(&b.state[2], op->rk_y));
  CU(cudaIpcGetMemHandle(&b.flags, op->p2p_flags));
  memset(blob, 0, RDYB_HALO_BLOB_BYTES);
  memcpy(blob, &b, sizeof(b));
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_halo_import(RdybOperator op, int32_t nranks, int32_t rank, const void *blobs) {
  RDYB_CHECK(op && blobs, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(op->p2p_flags, RDYB_ERR_USER, "rdyb_operator_halo_export must be called before rdyb_operator_halo_import");
  const RdybLayout &L = op->L;
  CU(cudaSetDevice(op->device));
  RDYB_CALL(adv_resolve(op));
  PeerTable &t = op->ptab;
  CU(rdyb_stream_sync(op->stream));
  for (int i = 0; i < op->p2p_nopened; ++i) cudaIpcCloseMemHandle(op->p2p_opened[i]); /* a repeated import replaces the earlier mapping */
  op->p2p_nopened = 0;
  op->p2p = false;
  CU(cudaMemsetAsync(op->p2p_flags, 0, sizeof(unsigned int) * (RDYB_MAX_PEERS + 8), op->stream));
  CU(rdyb_stream_sync(op->stream));
  memset(&t, 0, sizeof(t));
  t.num_peers = L.num_peers;
  { const char *env = getenv("RDYB_P2P_TIMEOUT_S"); if (env && atof(env) > 0.0) op->p2p_timeout_cycles = (long long)(atof(env) * 2.0e9); }
  if (!op->wait_cycles) RDYB_CALL(dalloc(&op->wait_cycles, 4));
  for (int p = 0; p < L.num_peers; ++p) {
    const int r = L.peer_ranks[p];
    RDYB_CHECK(r >= 0 && r < nranks && r != rank, RDYB_ERR_USER, "halo plan names rank %d (of %d) as a neighbour of rank %d", r, nranks, rank);
    HaloBlob b;
    memcpy(&b, (const char *)blobs + (size_t)r * RDYB_HALO_BLOB_BYTES, sizeof(b));
    RDYB_CHECK(b.magic == 0x52445942u && b.rank == r, RDYB_ERR_USER, "halo blob of rank %d is missing or corrupt", r);
    int q = -1;
    for (int j = 0; j < b.num_peers; ++j) if (b.peer_ranks[j] == rank) q = j;
    RDYB_CHECK(q >= 0, RDYB_ERR_USER, "rank %d sends to rank %d, which does not list it as a neighbour", rank, r);
    const int nsend = L.send_offsets[p + 1] - L.send_offsets[p], nrecv_there = b.recv_offsets[q + 1] - b.recv_offsets[q];
    RDYB_CHECK(nsend == nrecv_there, RDYB_ERR_USER, "rank %d sends %d cells to rank %d, which expects %d", rank, nsend, r, nrecv_there);
    void *ptr[4];
    const cudaIpcMemHandle_t hs[4] = {b.state[0], b.state[1], b.flags, b.state[2]};
    for (int k = 0; k < 4; ++k) {
      RDYB_CHECK(op->p2p_nopened < 4 * RDYB_MAX_PEERS, RDYB_ERR_SUP, "too many peer mappings");
      CU(cudaIpcOpenMemHandle(&ptr[k], hs[k], cudaIpcMemLazyEnablePeerAccess));
      op->p2p_opened[op->p2p_nopened++] = ptr[k];
    }
    t.state[p][0] = (double *)ptr[0]; t.state[p][1] = (double *)ptr[1]; t.state[p][2] = (double *)ptr[3];
    t.flag[p] = (unsigned int *)ptr[2] + q;
    t.stride[p] = b.stride; t.ghost0[p] = b.n_owned + b.recv_offsets[q];
    t.send_begin[p] = L.send_offsets[p];
  }
  t.send_begin[L.num_peers] = L.num_peers ? L.send_offsets[L.num_peers] : 0;
  op->p2p = L.num_peers > 0;
  op->p2p_epoch = 0;
  fpush_release(op);
  return RDYB_SUCCESS;
}

/* push the border cells of `state` (buffer index buf) into the neighbours' ghost slots and raise the step counter */
static int p2p_push(RdybOperator op, const double *state, int buf) {
  halo_push_kernel<<<1, 1024, 0, op->stream>>>(state, op->stride, op->send_cells, op->ptab, buf, ++op->p2p_epoch);
  LAUNCHED();
  CU(cudaGetLastError());
  return RDYB_SUCCESS;
}
/* send lists per frontier tile for the fused push: entry = {device cell, the neighbour's ghost slot in either buffer, its stride} */
static int p2p_fused_build(RdybOperator op) {
  if (op->fpush.built) return RDYB_SUCCESS;
  const RdybLayout &L = op->L;
  const auto &T = op->T;
  const int32_t nt = T.num_tiles, nit = T.num_interior_tiles, ns = L.send_offsets[L.num_peers];
  std::vector<int32_t> tile_of(ns), count(nt + 1, 0);
  for (int32_t i = 0; i < ns; ++i) {
    const int32_t c = L.send_cells[i];
    const int32_t t = (int32_t)(std::upper_bound(T.tile_cell_begin, T.tile_cell_begin + nt + 1, c) - T.tile_cell_begin) - 1;
    RDYB_CHECK(t >= nit && t < nt, RDYB_ERR_USER, "fused push: a send cell lies outside the frontier tiles");
    tile_of[i] = t;
    count[t + 1]++;
  }
  for (int32_t t = 0; t < nt; ++t) count[t + 1] += count[t];
  std::vector<int2> range(nt);
  unsigned int send_tiles = 0;
  for (int32_t t = 0; t < nt; ++t) { range[t] = make_int2(count[t], count[t + 1] - count[t]); send_tiles += range[t].y > 0; }
  std::vector<int32_t> src(ns), rs(ns), fill(count.begin(), count.end() - 1), order(nt);
  std::vector<double *> dst0(ns), dst1(ns);
  for (int p = 0; p < L.num_peers; ++p)
    for (int32_t i = L.send_offsets[p]; i < L.send_offsets[p + 1]; ++i) {
      const int32_t k = fill[tile_of[i]]++, g = op->ptab.ghost0[p] + (i - L.send_offsets[p]);
      src[k] = L.send_cells[i]; rs[k] = op->ptab.stride[p];
      dst0[k] = op->ptab.state[p][0] + g; dst1[k] = op->ptab.state[p][1] + g;
    }
  for (int32_t t = 0; t < nt; ++t) order[t] = t < nt - nit ? nit + t : t - (nt - nit); /* frontier tiles first */
  RDYB_CALL(upload(&op->fpush.range, range.data(), (size_t)nt));
  RDYB_CALL(upload(&op->fpush.src, src.data(), (size_t)ns));
  RDYB_CALL(upload(&op->fpush.rs, rs.data(), (size_t)ns));
  RDYB_CALL(upload(&op->fpush.order, order.data(), (size_t)nt));
  RDYB_CALL(upload(&op->fpush.dst[0], dst0.data(), (size_t)ns));
  RDYB_CALL(upload(&op->fpush.dst[1], dst1.data(), (size_t)ns));
  RDYB_CALL(dalloc(&op->fpush.counter, 1));
  CU(cudaMemset(op->fpush.counter, 0, sizeof(unsigned int)));
  op->fpush.send_tiles = send_tiles;
  op->fpush.launches = 0;
  op->fpush.built = true;
  return RDYB_SUCCESS;
}

static int p2p_wait(RdybOperator op) {
  halo_wait_kernel<<<1, 32, 0, op->stream>>>(op->p2p_flags, op->ptab.num_peers, op->p2p_epoch, op->p2p_timeout_cycles,
                                            reinterpret_cast<int *>(op->p2p_flags + RDYB_MAX_PEERS));
  LAUNCHED();
  CU(cudaGetLastError());
  return RDYB_SUCCESS;
}

/* RHS of a stage vector that lives in one of the three peer-mapped buffers (state[0], state[1], rk_y): border cells go straight into
 * the neighbours' ghost slots of the SAME buffer over NVLink, the interior cells run meanwhile, the frontier cells after the
 * neighbours' counters have arrived.  Safe while consecutive evaluations use different buffers: a rank starts the frontier of
 * evaluation k only after its neighbours have pushed their data of evaluation k, i.e. have finished evaluation k - 1, so nobody
 * is more than one evaluation ahead and a push never lands in a buffer its receiver still reads (step_rk4 orders its stages so). */
static int rhs_pass_p2p(RdybOperator op, double *in, int bufidx, double *out, double dt) {
  RDYB_CALL(p2p_push(op, in, bufidx));
  RDYB_CALL(launch_cells(op, in, out, dt, false, 0, op->NI, 0));
  RDYB_CALL(p2p_wait(op));
  RDYB_CALL(launch_cells(op, in, out, dt, false, op->NI, op->N, op->ncblocks_interior));
  return RDYB_SUCCESS;
}

/* one RHS (or fused Euler) evaluation of `in` into `out`, with the halo exchange of `in` overlapped with the interior cells */
static int rhs_pass(RdybOperator op, double *in, double *out, double dt, bool euler) {
  const bool exchange = op->L.num_peers > 0 && op->comm && !op->ghosts_from_caller;
  /* a partitioned mesh whose ghosts nobody fills would silently compute on stale values */
  RDYB_CHECK(op->L.num_peers == 0 || op->comm || op->ghosts_from_caller, RDYB_ERR_USER,
             "this rank has %d neighbour ranks but no communicator: call rdyb_operator_init_comm (or use rdyb_operator_apply_host, whose u_local carries the ghosts)", op->L.num_peers);
  if (op->bs.enabled) { /* CreateCeedEtaVerticesOperator runs before the flux operator in every RHS evaluation (src/operator.c:170-192) */
    bs2002_eta_vertex_kernel<<<grid_for(op->bs.nv, 128), 128, 0, op->stream>>>(in, op->bs.cz, op->stride, op->bs.vc_off, op->bs.vc_cell, op->bs.nv, op->cfg.tiny_h,
                                                                           op->bs.eta);
    LAUNCHED();
    CU(cudaGetLastError());
  }
  if (op->recon) {
    /* second order: state halo, gradients of the owned cells, gradient halo (two [3][stride] blocks: d/dx and d/dy, the two
     * passes of CommunicateCellGradients), then the reconstructed fluxes.  No overlap yet: this row is correctness-first. */
    RDYB_CHECK(op->L.num_peers == 0 || op->comm, RDYB_ERR_USER,
               "second order on a partitioned mesh exchanges cell gradients: call rdyb_operator_init_comm first");
    if (exchange) { RDYB_CALL(halo_begin(op, in)); RDYB_CALL(halo_end(op)); }
    SweArgs g;
    fill_args(op, g, in, out, dt);
    /* K4M in mode 2 derives every gradient it needs inside the tile kernel (single-rank layouts only): no gradient pass, no array */
    const bool in_kernel = op->kernel_kind >= 1 && op->has_tiled && !getenv("RDYB_MUSCL_K2M") && tiled_muscl_applicable(op->T, op->t_edge_off) &&
                           tiled_muscl_fixed_mode(op->T, op->t_edge_off, g, op->t_halo2 && op->t_halo_nbr) == 2;
    if (in_kernel) return launch_cells(op, in, out, dt, euler, 0, op->N, 0);
    muscl_gradient_kernel<<<grid_for(op->N, 128), 128, 0, op->stream>>>(g, op->grad, op->N);
    LAUNCHED();
    CU(cudaGetLastError());
    if (op->L.num_peers > 0) {
      RDYB_CALL(halo_begin(op, op->grad)); RDYB_CALL(halo_end(op));
      RDYB_CALL(halo_begin(op, op->grad + 3 * (size_t)op->stride)); RDYB_CALL(halo_end(op));
    }
    return launch_cells(op, in, out, dt, euler, 0, op->N, 0);
  }
  if (exchange) {
    RDYB_CALL(halo_begin(op, in));
    RDYB_CALL(launch_cells(op, in, out, dt, euler, 0, op->NI, 0));
    RDYB_CALL(halo_end(op));
    RDYB_CALL(launch_cells(op, in, out, dt, euler, op->NI, op->N, op->ncblocks_interior));
  } else {
    RDYB_CALL(launch_cells(op, in, out, dt, euler, 0, op->N, 0)); /* ghosts (if any) were filled by the caller */
  }
  return RDYB_SUCCESS;
}

/* ---------------- pipelined host-Vec apply ---------------- */
static void pipe_release(RdybOperator op) {
  auto &P = op->pipe;
  for (void *p : {(void *)P.order, (void *)P.local_to_dev, (void *)P.owned_to_dev, (void *)P.stage_in, (void *)P.stage_out})
    if (p) cudaFree(p);
  P.order = P.local_to_dev = P.owned_to_dev = nullptr; P.stage_in = P.stage_out = nullptr;
  P.built = P.usable = false;
}

/* launch order and readiness groups for the current tiled layout (host work, once per layout) */
static int pipe_build(RdybOperator op) {
  auto &P = op->pipe;
  pipe_release(op);
  P.built = true; P.tile_cells = op->T.tile_cells;
  const RdybLayout &L = op->L;
  const RdybTiledLayout &T = op->T;
  const int64_t nview = L.num_view_cells;
  const char *env = getenv("RDYB_HOST_PIPELINE"); /* "0" switches the pipeline off, a number sets the chunk count */
  int want = env ? atoi(env) : 16;
  const char *env_min = getenv("RDYB_HOST_PIPELINE_MIN_CELLS"); /* tests lower it to exercise the pipeline on small meshes */
  const int64_t min_cells = env_min ? atoll(env_min) : 262144;
  if (want <= 1 || nview < min_cells || op->N < min_cells || op->N < want) return RDYB_SUCCESS; /* small meshes: the copies are too short to overlap */
  if (want > 64) want = 64;
  P.nchunks = want;
  P.chunk_cells = (int32_t)((nview + want - 1) / want);
  P.out_cells = (op->N + want - 1) / want;
  const int nt = T.num_tiles, nit = T.num_interior_tiles;
  std::vector<int32_t> ready((size_t)nt, 0), tile_of((size_t)op->N, 0);
  for (int t = 0; t < nt; ++t) {
    const int32_t *d = T.tile_desc + 8 * (size_t)t;
    int r = 0;
    for (int32_t c = d[0]; c < d[0] + d[1]; ++c) { const int k = L.dev_to_local[c] / P.chunk_cells; if (k > r) r = k; tile_of[c] = t; }
    for (int32_t h = d[4]; h < d[4] + d[5]; ++h) { const int k = L.dev_to_local[T.halo_cells[h]] / P.chunk_cells; if (k > r) r = k; }
    ready[t] = r;
  }
  /* counting sort of the interior tiles, then of the frontier tiles, by the chunk that completes their inputs */
  std::vector<int32_t> order((size_t)nt);
  P.gi.assign((size_t)want + 1, 0); P.gf.assign((size_t)want + 1, 0);
  for (int t = 0; t < nit; ++t) ++P.gi[(size_t)ready[t] + 1];
  for (int t = nit; t < nt; ++t) ++P.gf[(size_t)ready[t] + 1];
  for (int c = 0; c < want; ++c) { P.gi[c + 1] += P.gi[c]; P.gf[c + 1] += P.gf[c]; }
  {
    std::vector<int32_t> ci(P.gi.begin(), P.gi.end() - 1), cf(P.gf.begin(), P.gf.end() - 1);
    for (int t = 0; t < nit; ++t) order[(size_t)ci[ready[t]]++] = t;
    for (int t = nit; t < nt; ++t) order[(size_t)nit + cf[ready[t]]++] = t;
  }
  for (int c = 0; c <= want; ++c) P.gf[c] += nit; /* absolute launch positions */
  const int nout = (op->N + P.out_cells - 1) / P.out_cells;
  P.out_need.assign((size_t)nout, 0);
  for (int32_t o = 0; o < op->N; ++o) {
    const int r = ready[tile_of[L.owned_to_dev[o]]];
    int32_t &need = P.out_need[(size_t)(o / P.out_cells)];
    if (r > need) need = r;
  }
  RDYB_CALL(upload(&P.order, order.data(), (size_t)nt));
  RDYB_CALL(upload(&P.local_to_dev, L.local_to_dev, (size_t)nview));
  RDYB_CALL(upload(&P.owned_to_dev, L.owned_to_dev, (size_t)op->N));
  RDYB_CALL(dalloc(&P.stage_in, 3 * (size_t)nview));
  RDYB_CALL(dalloc(&P.stage_out, 3 * (size_t)op->N));
  if (!P.s_in) { CU(cudaStreamCreateWithFlags(&P.s_in, cudaStreamNonBlocking)); CU(cudaStreamCreateWithFlags(&P.s_out, cudaStreamNonBlocking)); }
  while ((int)P.ev_in.size() < want) { cudaEvent_t e; CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); P.ev_in.push_back(e); }
  while ((int)P.ev_out.size() < nout) { cudaEvent_t e; CU(cudaEventCreateWithFlags(&e, cudaEventDisableTiming)); P.ev_out.push_back(e); }
  P.usable = true;
  return RDYB_SUCCESS;
}

static int apply_host_pipelined(RdybOperator op, double dt, const double *u_local, double *f_global) {
  auto &P = op->pipe;
  const int64_t nview = op->L.num_view_cells;
  double *state = op->state[op->cur];
  const int nout = (int)P.out_need.size();
  int kout = 0;
  ncclComm_t saved = op->comm;
  op->comm = nullptr; /* ghosts come from the caller */
  int ierr = RDYB_SUCCESS;
  auto body = [&]() -> int {
    for (int c = 0; c < P.nchunks; ++c) {
      const int64_t c0 = (int64_t)c * P.chunk_cells;
      const int32_t n = (int32_t)std::min<int64_t>(P.chunk_cells, nview - c0);
      if (n > 0) {
        CU(cudaMemcpyAsync(P.stage_in + 3 * c0, u_local + 3 * c0, sizeof(double) * 3 * (size_t)n, cudaMemcpyHostToDevice, P.s_in));
        CU(cudaEventRecord(P.ev_in[c], P.s_in));
        CU(cudaStreamWaitEvent(op->stream, P.ev_in[c], 0));
        aos_scatter_soa_kernel<<<grid_for(n, 256), 256, 0, op->stream>>>(P.stage_in + 3 * c0, P.local_to_dev + c0, state, n, op->stride);
        LAUNCHED();
        CU(cudaGetLastError());
      }
      op->td.order = P.order;
      if (P.gi[c + 1] > P.gi[c]) { op->td.sub_begin = P.gi[c]; op->td.sub_end = P.gi[c + 1]; RDYB_CALL(launch_cells(op, state, op->F, dt, false, 0, op->NI, 0)); }
      if (P.gf[c + 1] > P.gf[c]) { op->td.sub_begin = P.gf[c]; op->td.sub_end = P.gf[c + 1]; RDYB_CALL(launch_cells(op, state, op->F, dt, false, op->NI, op->N, op->ncblocks_interior)); }
      for (; kout < nout && P.out_need[kout] <= c; ++kout) {
        const int64_t o0 = (int64_t)kout * P.out_cells;
        const int32_t m = (int32_t)std::min<int64_t>(P.out_cells, op->N - o0);
        soa_gather_aos_kernel<<<grid_for(m, 256), 256, 0, op->stream>>>(op->F, P.owned_to_dev + o0, P.stage_out + 3 * o0, m, op->stride);
        LAUNCHED();
        CU(cudaGetLastError());
        CU(cudaEventRecord(P.ev_out[kout], op->stream));
        CU(cudaStreamWaitEvent(P.s_out, P.ev_out[kout], 0));
        CU(cudaMemcpyAsync(f_global + 3 * o0, P.stage_out + 3 * o0, sizeof(double) * 3 * (size_t)m, cudaMemcpyDeviceToHost, P.s_out));
      }
    }
    return RDYB_SUCCESS;
  };
  ierr = body();
  op->td.order = nullptr; op->td.sub_begin = op->td.sub_end = 0;
  op->comm = saved;
  rdyb_stream_sync(P.s_in);
  rdyb_stream_sync(op->stream);
  cudaError_t e = rdyb_stream_sync(P.s_out);
  RDYB_CALL(ierr);
  if (e != cudaSuccess) return rdyb_fail(RDYB_ERR_CUDA, "CUDA error in the pipelined host-Vec apply (%s)", cudaGetErrorString(e));
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_apply_host(RdybOperator op, double dt, const double *u_local, double *f_global) {
  RDYB_CHECK(op && u_local && f_global, RDYB_ERR_ARG, "NULL argument");
  CU(cudaSetDevice(op->device));
  RDYB_CALL(adv_resolve(op));
  if ((op->kernel_kind == 2 || op->kernel_kind == 3) && op->has_tiled && !op->recon) { /* K3 / K4: chunked, overlapped copies */
    if (!op->pipe.built || op->pipe.tile_cells != op->T.tile_cells) RDYB_CALL(pipe_build(op));
    if (op->pipe.usable) return apply_host_pipelined(op, dt, u_local, f_global);
  }
  /* u_local (AoS, view order, ghosts included) -> device SoA */
  CU(cudaMemcpyAsync(op->stage, u_local, sizeof(double) * 3 * (size_t)op->L.num_view_cells, cudaMemcpyHostToDevice, op->stream));
  aos_to_soa_kernel<<<grid_for(op->NL, 256), 256, 0, op->stream>>>(op->stage, op->dev_to_local, op->state[op->cur], op->NL, op->stride);
  LAUNCHED();
  CU(cudaGetLastError());
  /* ghosts come from the caller here (DMGlobalToLocal already ran): no state exchange */
  op->ghosts_from_caller = true;
  int ierr = rhs_pass(op, op->state[op->cur], op->F, dt, false);
  op->ghosts_from_caller = false;
  RDYB_CALL(ierr);
  soa_to_aos_kernel<<<grid_for(op->N, 256), 256, 0, op->stream>>>(op->F, op->dev_to_owned, op->stage, op->N, op->stride);
  LAUNCHED();
  CU(cudaGetLastError());
  CU(cudaMemcpyAsync(f_global, op->stage, sizeof(double) * 3 * op->N, cudaMemcpyDeviceToHost, op->stream));
  CU(rdyb_stream_sync(op->stream));
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_apply(RdybOperator op, double dt, double *f_global) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "NULL operator");
  CU(cudaSetDevice(op->device));
  RDYB_CALL(adv_resolve(op));
  RDYB_CALL(rhs_pass(op, op->state[op->cur], op->F, dt, false));
  if (f_global) {
    soa_to_aos_kernel<<<grid_for(op->N, 256), 256, 0, op->stream>>>(op->F, op->dev_to_owned, op->stage, op->N, op->stride);
    LAUNCHED();
    CU(cudaGetLastError());
    CU(cudaMemcpyAsync(f_global, op->stage, sizeof(double) * 3 * op->N, cudaMemcpyDeviceToHost, op->stream));
    CU(rdyb_stream_sync(op->stream));
  }
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_step_euler(RdybOperator op, double dt, int32_t nsteps) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "NULL operator");
  RDYB_CHECK(nsteps >= 0, RDYB_ERR_ARG, "negative step count");
  CU(cudaSetDevice(op->device));
  RDYB_CALL(adv_resolve(op));
  if (op->p2p && nsteps > 0 && !op->recon) {
    /* peer-memory halo: the two state buffers alternate, so a neighbour's push for step s+1 lands in the buffer this rank
     * does not read during step s, and its counter (raised after its frontier tiles) orders everything else.
     * Epoch accounting is the same on every path, so that ranks which pick different paths (a rank-local decision: tile shapes,
     * momentum sources) stay in step: one priming push per call (+1), then +1 after every step, the last one included. */
    RDYB_CALL(p2p_push(op, op->state[op->cur], op->cur));
    /* the timed region of rdyb_operator_last_kernel_ms starts on the device once every neighbour has primed too: host launch
     * skew between the ranks stays outside it */
    RDYB_CALL(p2p_wait(op));
    CU(cudaEventRecord(op->ev_t0, op->stream));
    /* K4 over the whole range: the wait for the neighbours' border cells happens inside the kernel, before a CTA's first frontier tile */
    SweArgs probe;
    fill_args(op, probe, op->state[0], op->state[1], dt);
    const bool one_launch = op->has_tiled && op->kernel_kind == 3 && op->NI < op->N && !getenv("RDYB_P2P_SPLIT_LAUNCH") &&
                            tiled_fixed_applicable(op->T.tile_cells, op->T.max_halo, op->T.max_edges, probe);
    /* fused push: the kernel of step s also delivers the border cells of step s+1 (and raises epoch + 1), so the loop has no
     * push kernel at all */
    const bool fused = one_launch && op->L.send_offsets[op->L.num_peers] > 0 && !getenv("RDYB_P2P_NO_FUSED_PUSH");
    if (fused) RDYB_CALL(p2p_fused_build(op));
    for (int32_t s = 0; s < nsteps; ++s) {
      double *in = op->state[op->cur], *out = op->state[op->cur ^ 1];
      if (one_launch) {
        op->wait_in_kernel = true;
        op->fpush.active = fused;
        const int rc = launch_cells(op, in, out, dt, true, 0, op->N, 0);
        op->wait_in_kernel = false;
        op->fpush.active = false;
        RDYB_CALL(rc);
        op->cur ^= 1;
        if (fused) ++op->p2p_epoch; /* raised in the neighbours by the kernel itself */
        else RDYB_CALL(p2p_push(op, op->state[op->cur], op->cur));
        continue;
      }
      RDYB_CALL(launch_cells(op, in, out, dt, true, 0, op->NI, 0));                       /* interior cells: no ghost needed */
      RDYB_CALL(p2p_wait(op));                                                          /* neighbours' border cells have landed */
      RDYB_CALL(launch_cells(op, in, out, dt, true, op->NI, op->N, op->ncblocks_interior)); /* frontier cells */
      op->cur ^= 1;
      RDYB_CALL(p2p_push(op, op->state[op->cur], op->cur));
    }
  } else {
    CU(cudaEventRecord(op->ev_t0, op->stream));
    for (int32_t s = 0; s < nsteps; ++s) {
      RDYB_CALL(rhs_pass(op, op->state[op->cur], op->state[op->cur ^ 1], dt, true));
      op->cur ^= 1;
    }
  }
  CU(cudaEventRecord(op->ev_t1, op->stream));
  op->last_launches = nsteps;
  return RDYB_SUCCESS;
}

/* ticks (clock64, ~SM clock) the in-kernel neighbour waits of this rank took since the last call: sum over the waiting CTAs, the
 * longest single wait and the number of waits (one per CTA that owns a frontier tile, per launch); resets the counters (syncs) */
extern "C" int rdyb_operator_get_wait_cycles(RdybOperator op, uint64_t *sum, uint64_t *longest, uint64_t *waits) {
  RDYB_CHECK(op && sum && longest && waits, RDYB_ERR_ARG, "NULL argument");
  *sum = 0; *longest = 0; *waits = 0;
  if (!op->wait_cycles) return RDYB_SUCCESS;
  CU(cudaSetDevice(op->device));
  unsigned long long h[4] = {0, 0, 0, 0};
  CU(cudaMemcpyAsync(h, op->wait_cycles, sizeof(h), cudaMemcpyDeviceToHost, op->stream));
  CU(cudaMemsetAsync(op->wait_cycles, 0, sizeof(h), op->stream));
  CU(rdyb_stream_sync(op->stream));
  *sum = h[0]; *longest = h[1]; *waits = h[2];
  return RDYB_SUCCESS;
}


/* ---------------- Bradford-Sanders (2002) well balancing, libCEED semantics ----------------
 * `physics.flow.well_balancing: bradford_sanders_2002` changes the reference's libCEED operator only (src/operator.c:170-192): a
 * vertex pass computes the water-surface elevation at every mesh vertex from the cells around it (SWEEtaVertex), and every edge flux
 * gains g dhv^2 / 24 (cn, sn), dhv = the change of the vertex depth along the edge (ComputeDhv).  The PETSc operator this library
 * replaces ignores the option (src/operator.c:202-209), and so does rdyb_operator_create: BS2002 is switched on by THIS call, which
 * hands over the vertex data the mesh view does not carry.  Triangles, serial meshes, first order, K1. */
extern "C" int rdyb_operator_enable_bs2002(RdybOperator op, const RdybVertexView *vv) {
  RDYB_CHECK(op && vv && vv->vertex_z && vv->cell_vertex_offsets && vv->cell_vertex_ids && vv->edge_vertex_ids, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(op->L.num_peers == 0, RDYB_ERR_SUP, "BS2002 well balancing is implemented for serial meshes");
  RDYB_CHECK(!op->recon && op->cfg.well_balancing != RDYB_WELL_BALANCING_HR, RDYB_ERR_USER, "BS2002 well balancing goes with the first-order Roe flux");
  CU(cudaSetDevice(op->device));
  const RdybLayout &L = op->L;
  const int32_t nv = vv->num_vertices, nc = op->NL;
  RDYB_CHECK(nv > 0, RDYB_ERR_ARG, "no vertices");
  std::vector<double> cz(3 * (size_t)op->stride, 0.0);
  std::vector<int32_t> off((size_t)nv + 1, 0);
  for (int32_t c = 0; c < nc; ++c) { /* c = view-local cell id */
    const int32_t o = vv->cell_vertex_offsets[c], n = vv->cell_vertex_offsets[c + 1] - o;
    RDYB_CHECK(n == 3, RDYB_ERR_SUP, "BS2002 well balancing is defined on triangles (cell %d has %d vertices), src/swe/swe_well_balancing.h:30", c, n);
    double z[3];
    for (int k = 0; k < 3; ++k) {
      const int32_t v = vv->cell_vertex_ids[o + k];
      RDYB_CHECK(v >= 0 && v < nv, RDYB_ERR_ARG, "vertex id %d out of range", v);
      z[k] = vv->vertex_z[v];
      off[(size_t)v + 1]++;
    }
    std::sort(z, z + 3);
    const int32_t d = L.local_to_dev[c];
    for (int k = 0; k < 3; ++k) cz[(size_t)k * op->stride + d] = z[k];
  }
  for (int32_t v = 0; v < nv; ++v) off[(size_t)v + 1] += off[v];
  std::vector<int32_t> cells((size_t)off[nv]), fill(off.begin(), off.end() - 1);
  for (int32_t c = 0; c < nc; ++c) /* ascending view-local cell id per vertex: the summation order of the oracle */
    for (int k = 0; k < 3; ++k) cells[(size_t)fill[vv->cell_vertex_ids[vv->cell_vertex_offsets[c] + k]]++] = L.local_to_dev[c];
  const size_t W = (size_t)L.width * (size_t)L.stride;
  std::vector<int32_t> vb(W, 0), ve(W, 0);
  for (size_t q = 0; q < W; ++q) {
    const int32_t e = L.ell_edge[q];
    if (e >= 0) { vb[q] = vv->edge_vertex_ids[2 * (size_t)e]; ve[q] = vv->edge_vertex_ids[2 * (size_t)e + 1]; }
  }
  auto &B = op->bs;
  for (void *q : {(void *)B.vz, (void *)B.eta, (void *)B.cz, (void *)B.vc_off, (void *)B.vc_cell, (void *)B.vb, (void *)B.ve})
    if (q) cudaFree(q);
  B = {};
  RDYB_CALL(upload(&B.vz, vv->vertex_z, (size_t)nv));
  RDYB_CALL(upload(&B.eta, vv->vertex_z, (size_t)nv)); /* CreateCeedEtaVerticesVector: the vertex elevations until the first pass */
  RDYB_CALL(upload(&B.cz, cz.data(), cz.size()));
  RDYB_CALL(upload(&B.vc_off, off.data(), off.size()));
  RDYB_CALL(upload(&B.vc_cell, cells.data(), cells.size()));
  RDYB_CALL(upload(&B.vb, vb.data(), W));
  RDYB_CALL(upload(&B.ve, ve.data(), W));
  B.nv = nv;
  B.enabled = true;
  adv_release_graph(op);
  return RDYB_SUCCESS;
}

/* the vertex water-surface elevations of the last right-hand-side evaluation (tests) */
extern "C" int rdyb_operator_get_eta_vertices(RdybOperator op, int32_t num_vertices, double *eta) {
  RDYB_CHECK(op && eta, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(op->bs.enabled, RDYB_ERR_USER, "BS2002 well balancing is not enabled");
  RDYB_CHECK(num_vertices == op->bs.nv, RDYB_ERR_ARG_SIZ, "Wrong number of vertices: %d (should be %d)", num_vertices, op->bs.nv);
  CU(cudaSetDevice(op->device));
  CU(cudaMemcpyAsync(eta, op->bs.eta, sizeof(double) * (size_t)num_vertices, cudaMemcpyDeviceToHost, op->stream));
  CU(rdyb_stream_sync(op->stream));
  return RDYB_SUCCESS;
}

/* ---------------- device-side time loop: one coupling interval without a host round trip ----------------
 * RDyAdvance (src/rdyadvance.c:261-383) = [adaptive-dt rule from the last Courant number, :303-343] + TSSolve to the next
 * coupling time with TS_EXACTFINALTIME_MATCHSTEP + UpdateOperatorDiagnostics (:366-372).  The reference (and round 1 of this
 * library) needs the Courant number on the HOST to pick dt and the step count - one synchronisation + reduction per interval, a
 * round trip per step when the coupling interval is one step.  Here the whole interval is ONE CUDA graph launch:
 *   adv_begin_kernel   reduces the Courant slots the last step left behind, applies the dt rule, splits the interval into nfull
 *                      steps of dt + one shortened step (the same arithmetic as the host path, so both give the same bits), and
 *                      arms the loop;
 *   a WHILE conditional node (cudaGraphCondTypeWhile) whose body is two fused Euler steps, state buffer 0 -> 1 and 1 -> 0; each
 *                      step kernel reads its dt and its turn from the control block (SweArgs::ctl) and returns at once when it is
 *                      not its turn; adv_step_kernel after each advances the step counter, picks the next dt, clears the Courant
 *                      slots before the LAST step (the reference reports the Courant number of the last RHS evaluation only,
 *                      src/rdysetup.c:1136) and sets the loop condition with cudaGraphSetConditional.
 * The host learns t, dt, the step count or which state buffer is current only when a getter asks (adv_resolve). */
__global__ void __launch_bounds__(256) adv_begin_kernel(AdvCtl *ctl, double *cmax, long long *carg, int32_t nslots, double interval, int adaptive,
                                                         double target, double maxinc, cudaGraphConditionalHandle handle) {
  __shared__ double red[256];
  double m = 0.0;
  for (int i = threadIdx.x; i < nslots; i += 256) m = fmax(m, cmax[i]);
  red[threadIdx.x] = m;
  __syncthreads();
  for (int off = 128; off > 0; off >>= 1) {
    if (threadIdx.x < off) red[threadIdx.x] = fmax(red[threadIdx.x], red[threadIdx.x + off]);
    __syncthreads();
  }
  __shared__ int total_s;
  if (threadIdx.x == 0) {
    const double courant = red[0];
    if (ctl->steps > 0) { ctl->courant = courant; ctl->courant_valid = courant > 0.0; } /* UpdateOperatorDiagnostics of the previous interval */
    double dt = ctl->dt;
    if (adaptive && ctl->courant_valid) { /* src/rdyadvance.c:303-343 */
      if (ctl->courant < target) {
        const double factor = fmin(__ddiv_rn(target, ctl->courant), maxinc);
        dt = __dmul_rn(dt, factor);
        if (dt > interval) dt = interval;
      } else {
        dt = __dmul_rn(dt, __ddiv_rn(target, ctl->courant));
      }
    }
    ctl->dt = dt;
    ctl->t_end = ctl->t + interval;
    const double span = ctl->t_end - ctl->t; /* as the host path forms it (rdy_driver.c): the same bits, hence the same step split */
    /* no FMA contraction here: the host path (gcc, rdy_driver.c) rounds every operation, and one ulp in `rest` is one ulp in the
     * last step's dt */
    int nfull = (int)floor(__dadd_rn(__dmul_rn(__ddiv_rn(span, dt), 1.0 + 1e-14), 1e-12));
    if (nfull < 0) nfull = 0;
    double rest = __dsub_rn(span, __dmul_rn((double)nfull, dt));
    if (rest < __dmul_rn(1e-12, fmax(span, dt))) rest = 0.0;
    ctl->nfull = nfull; ctl->rest = rest; ctl->k = 0;
    ctl->total = nfull + (rest > 0.0 ? 1 : 0);
    ctl->dt_step = nfull > 0 ? dt : rest;
    ctl->active = ctl->total > 0;
    total_s = ctl->total;
    cudaGraphSetConditional(handle, ctl->active ? 1u : 0u);
  }
  __syncthreads();
  if (total_s == 1) /* the first step is also the last: its Courant number is the one to report */
    for (int i = threadIdx.x; i < nslots; i += 256) { cmax[i] = 0.0; carg[i] = -1; }
}

__global__ void __launch_bounds__(256) adv_step_kernel(AdvCtl *ctl, int32_t buf, double *cmax, long long *carg, int32_t nslots,
                                                        cudaGraphConditionalHandle handle) {
  __shared__ int clear_s;
  if (threadIdx.x == 0) {
    clear_s = 0;
    if (ctl->active && ctl->cur == buf) { /* the step kernel before this one ran */
      ctl->cur = buf ^ 1;
      ctl->steps += 1;
      const int k = ++ctl->k;
      if (k >= ctl->total) {
        ctl->active = 0;
        ctl->t = ctl->t_end; /* TS_EXACTFINALTIME_MATCHSTEP: the interval ends exactly on the coupling time */
      } else {
        ctl->t = __dadd_rn(ctl->t, ctl->dt_step);
        ctl->dt_step = k < ctl->nfull ? ctl->dt : ctl->rest;
        clear_s = (k == ctl->total - 1);
      }
    }
    cudaGraphSetConditional(handle, ctl->active ? 1u : 0u);
  }
  __syncthreads();
  if (clear_s)
    for (int i = threadIdx.x; i < nslots; i += 256) { cmax[i] = 0.0; carg[i] = -1; }
}

static void adv_release_graph(RdybOperator op) {
  if (op->adv.exec) cudaGraphExecDestroy(op->adv.exec);
  if (op->adv.graph) cudaGraphDestroy(op->adv.graph);
  op->adv.exec = nullptr; op->adv.graph = nullptr;
}

/* bring the host's view (which state buffer is current) up to date after device-side intervals; the only place that waits */
static int adv_resolve(RdybOperator op) {
  if (!op->adv.dirty) return RDYB_SUCCESS;
  CU(rdyb_event_sync(op->adv.ev_mirror));
  op->cur = op->adv.mirror->cur;
  op->adv.dirty = false;
  return RDYB_SUCCESS;
}

static bool adv_kernel_supported(RdybOperator op, double dt_probe) {
  if (op->recon || op->L.num_peers > 0 || op->bs.enabled) return false; /* second order / partitioned meshes / BS2002 (a vertex pass per step): host-driven path */
  if (op->kernel_kind == 0) return true;            /* K1 */
  if (op->kernel_kind == 3 && op->has_tiled) {      /* K4, when the mesh fits its compiled shapes and options */
    SweArgs probe;
    fill_args(op, probe, op->state[0], op->state[1], dt_probe);
    return tiled_fixed_applicable(op->T.tile_cells, op->T.max_halo, op->T.max_edges, probe);
  }
  return false;
}

extern "C" int rdyb_operator_time_set(RdybOperator op, double t, double dt) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "NULL operator");
  RDYB_CHECK(dt > 0.0, RDYB_ERR_USER, "time step must be positive");
  CU(cudaSetDevice(op->device));
  RDYB_CALL(adv_resolve(op));
  if (!op->adv.ctl) {
    RDYB_CALL(dalloc(&op->adv.ctl, 1));
    CU(cudaMallocHost((void **)&op->adv.mirror, sizeof(AdvCtl)));
    CU(cudaEventCreateWithFlags(&op->adv.ev_mirror, cudaEventDisableTiming));
  }
  AdvCtl c;
  memset(&c, 0, sizeof(c));
  c.t = t; c.t_end = t; c.dt = dt; c.dt_step = dt; c.cur = op->cur;
  *op->adv.mirror = c;
  CU(cudaMemcpyAsync(op->adv.ctl, op->adv.mirror, sizeof(AdvCtl), cudaMemcpyHostToDevice, op->stream));
  CU(rdyb_stream_sync(op->stream));
  return RDYB_SUCCESS;
}

/* t, dt of the next interval's first guess, steps taken, Courant number of the last finished interval's last step (the value the
 * dt rule will use); waits for the queued intervals */
extern "C" int rdyb_operator_time_get(RdybOperator op, double *t, double *dt, int64_t *steps, double *courant) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "NULL operator");
  RDYB_CHECK(op->adv.ctl, RDYB_ERR_USER, "rdyb_operator_time_set has not been called");
  CU(cudaSetDevice(op->device));
  CU(cudaMemcpyAsync(op->adv.mirror, op->adv.ctl, sizeof(AdvCtl), cudaMemcpyDeviceToHost, op->stream));
  CU(cudaEventRecord(op->adv.ev_mirror, op->stream));
  op->adv.dirty = true;
  RDYB_CALL(adv_resolve(op));
  RDYB_CALL(p2p_check_status(op));
  if (t) *t = op->adv.mirror->t;
  if (dt) *dt = op->adv.mirror->dt;
  if (steps) *steps = op->adv.mirror->steps;
  if (courant) *courant = op->adv.mirror->courant;
  return RDYB_SUCCESS;
}

/* returns RDYB_ERR_SUP (and does nothing) when this operator's configuration has no device-side loop: the caller falls back to
 * rdyb_operator_step_euler with host-side bookkeeping */
extern "C" int rdyb_operator_advance_device(RdybOperator op, double interval, int adaptive, double target_courant, double max_increase_factor) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "NULL operator");
  RDYB_CHECK(interval > 0.0, RDYB_ERR_USER, "coupling interval must be positive");
  RDYB_CHECK(op->adv.ctl, RDYB_ERR_USER, "rdyb_operator_time_set has not been called");
  CU(cudaSetDevice(op->device));
  if (!adv_kernel_supported(op, 1.0) || getenv("RDYB_NO_DEVICE_LOOP")) return RDYB_ERR_SUP;
  const int opts = (op->opt_courant ? 1 : 0) | (op->opt_fdiv ? 2 : 0) | (op->opt_prim ? 4 : 0) | (op->opt_bflux ? 8 : 0) | (op->mom_src ? 16 : 0);
  auto &A = op->adv;
  const bool stale = !A.exec || A.interval != interval || A.adaptive != adaptive || A.target != target_courant || A.maxinc != max_increase_factor ||
                     A.kernel_kind != op->kernel_kind || A.tile_cells != (op->has_tiled ? op->T.tile_cells : 0) || A.opts != opts || A.max_ctas != op->max_ctas;
  if (stale) {
    adv_release_graph(op);
    const bool courant_was = op->opt_courant;
    op->opt_courant = true; /* the dt rule needs it */
    /* configure the kernels outside the capture (function attributes, occupancy): both directions once, gated off */
    AdvCtl off;
    memset(&off, 0, sizeof(off));
    off.cur = -1;
    AdvCtl *gate_off = nullptr;
    RDYB_CALL(dalloc(&gate_off, 1));
    CU(cudaMemcpyAsync(gate_off, &off, sizeof(off), cudaMemcpyHostToDevice, op->stream));
    auto launch_dir = [&](int buf, const AdvCtl *ctl) -> int {
      op->adv_ctl_arg = ctl; op->adv_buf_arg = buf;
      const int rc = launch_cells(op, op->state[buf], op->state[buf ^ 1], 0.0, true, 0, op->N, 0);
      op->adv_ctl_arg = nullptr;
      return rc;
    };
    int rc = launch_dir(0, gate_off);
    if (!rc) rc = launch_dir(1, gate_off);
    CU(rdyb_stream_sync(op->stream));
    cudaFree(gate_off);
    (void)courant_was;
    if (rc) return rc;
    cudaGraph_t g = nullptr;
    CU(cudaGraphCreate(&g, 0));
    A.graph = g;
    cudaGraphConditionalHandle handle;
    CU(cudaGraphConditionalHandleCreate(&handle, g, 0, 0));
    /* node 1: the dt rule */
    CU(cudaStreamBeginCaptureToGraph(op->stream, g, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
    adv_begin_kernel<<<1, 256, 0, op->stream>>>(A.ctl, op->cmax, op->carg, op->ncblocks, interval, adaptive, target_courant, max_increase_factor, handle);
    cudaGraph_t same = nullptr;
    CU(cudaStreamEndCapture(op->stream, &same));
    size_t nnodes = 0;
    CU(cudaGraphGetNodes(g, nullptr, &nnodes));
    RDYB_CHECK(nnodes == 1, RDYB_ERR_CUDA, "device-side loop: unexpected graph shape (%zu nodes)", nnodes);
    cudaGraphNode_t begin_node;
    CU(cudaGraphGetNodes(g, &begin_node, &nnodes));
    /* node 2: while (active) { step 0->1; bookkeeping; step 1->0; bookkeeping } */
    cudaGraphNodeParams &cp = *static_cast<cudaGraphNodeParams *>(calloc(1, sizeof(cudaGraphNodeParams))); /* (its default constructor is deleted; a few bytes per graph build) */
    cp.type = cudaGraphNodeTypeConditional;
    cp.conditional.handle = handle;
    cp.conditional.type = cudaGraphCondTypeWhile;
    cp.conditional.size = 1;
    cudaGraphNode_t cond_node;
    CU(cudaGraphAddNode(&cond_node, g, &begin_node, 1, &cp));
    cudaGraph_t body = cp.conditional.phGraph_out[0];
    CU(cudaStreamBeginCaptureToGraph(op->stream, body, nullptr, nullptr, 0, cudaStreamCaptureModeRelaxed));
    rc = RDYB_SUCCESS;
    for (int buf = 0; buf < 2 && !rc; ++buf) {
      rc = launch_dir(buf, A.ctl);
      adv_step_kernel<<<1, 256, 0, op->stream>>>(A.ctl, buf, op->cmax, op->carg, op->ncblocks, handle);
    }
    cudaGraph_t body_same = nullptr;
    cudaError_t ce = cudaStreamEndCapture(op->stream, &body_same);
    if (rc) { adv_release_graph(op); return rc; }
    if (ce != cudaSuccess) { adv_release_graph(op); return rdyb_fail(RDYB_ERR_CUDA, "device-side loop: capture failed (%s)", cudaGetErrorString(ce)); }
    ce = cudaGraphInstantiate(&A.exec, g, 0);
    if (ce != cudaSuccess) { adv_release_graph(op); return rdyb_fail(RDYB_ERR_CUDA, "device-side loop: graph instantiation failed (%s)", cudaGetErrorString(ce)); }
    A.interval = interval; A.adaptive = adaptive; A.target = target_courant; A.maxinc = max_increase_factor;
    A.kernel_kind = op->kernel_kind; A.tile_cells = op->has_tiled ? op->T.tile_cells : 0; A.opts = opts; A.max_ctas = op->max_ctas;
  }
  CU(cudaGraphLaunch(A.exec, op->stream));
  LAUNCHED();
  /* keep the pinned mirror fresh without waiting: a getter only has to wait for this copy */
  CU(cudaMemcpyAsync(A.mirror, A.ctl, sizeof(AdvCtl), cudaMemcpyDeviceToHost, op->stream));
  CU(cudaEventRecord(A.ev_mirror, op->stream));
  A.dirty = true;
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_reset_diagnostics(RdybOperator op);
extern "C" int rdyb_operator_step_rk4(RdybOperator op, double dt, int32_t nsteps) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "NULL operator");
  CU(cudaSetDevice(op->device));
  RDYB_CALL(adv_resolve(op));
  const size_t S = (size_t)op->stride;
  for (int i = 0; i < 4; ++i)
    if (!op->rk_k[i]) RDYB_CALL(dalloc(&op->rk_k[i], 3 * S));
  if (!op->rk_y) RDYB_CALL(dalloc(&op->rk_y, 3 * S));
  /* classical tableau, as PETSc TSRK4: Y_i = U + dt*a_i*K_{i-1}; U += dt*(K0/6 + K1/3 + K2/3 + K3/6) */
  const double a[4] = {0.0, 0.5, 0.5, 1.0};
  const double bw[4] = {1.0 / 6.0, 1.0 / 3.0, 1.0 / 3.0, 1.0 / 6.0};
  const int g = grid_for(op->N, 256);
  CU(cudaEventRecord(op->ev_t0, op->stream));
  /* the stage vectors alternate between the spare state buffer and rk_y, so that consecutive RHS evaluations never read the same
   * buffer: U, S, R, S | U, S, R, S ... - what the peer-memory halo needs (rhs_pass_p2p); harmless on one GPU */
  const bool p2p = op->p2p && !op->recon && !getenv("RDYB_RK4_NCCL");
  for (int32_t s = 0; s < nsteps; ++s) {
    double *U = op->state[op->cur];
    for (int i = 0; i < 4; ++i) {
      double *Y = U;
      int ybuf = op->cur;
      if (i > 0) {
        Y = (i & 1) ? op->state[op->cur ^ 1] : op->rk_y;
        ybuf = (i & 1) ? (op->cur ^ 1) : 2;
        stage_axpy_kernel<<<g, 256, 0, op->stream>>>(U, op->rk_k[i - 1], dt * a[i], Y, op->N, op->stride);
        LAUNCHED();
        CU(cudaGetLastError());
      }
      /* OperatorRHSFunction resets the diagnostics in every call (src/rdysetup.c:1136), so after TSSolve they hold the LAST stage's
       * Courant number only: clear the slots before the final stage of the final step */
      if (i == 3 && s == nsteps - 1 && op->opt_courant) RDYB_CALL(rdyb_operator_reset_diagnostics(op));
      if (p2p) RDYB_CALL(rhs_pass_p2p(op, Y, ybuf, op->rk_k[i], dt));
      else RDYB_CALL(rhs_pass(op, Y, op->rk_k[i], dt, false));
    }
    rk4_combine_kernel<<<g, 256, 0, op->stream>>>(U, op->rk_k[0], op->rk_k[1], op->rk_k[2], op->rk_k[3], dt * bw[0], dt * bw[1], dt * bw[2],
                                                  dt * bw[3], op->N, op->stride);
    LAUNCHED();
    CU(cudaGetLastError());
  }
  CU(cudaEventRecord(op->ev_t1, op->stream));
  op->last_launches = nsteps; /* rdyb_operator_last_kernel_ms: ms per STEP (4 RHS evaluations + stage kernels) */
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_synchronize(RdybOperator op) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "NULL operator");
  CU(cudaSetDevice(op->device));
  CU(rdyb_stream_sync(op->comm_stream));
  CU(rdyb_stream_sync(op->stream));
  return p2p_check_status(op);
}

/* ---------------- diagnostics ---------------- */
extern "C" int rdyb_operator_reset_diagnostics(RdybOperator op) {
  RDYB_CHECK(op, RDYB_ERR_ARG, "NULL operator");
  CU(cudaSetDevice(op->device));
  CU(cudaMemsetAsync(op->cmax, 0, sizeof(double) * (size_t)op->ncblocks, op->stream));
  CU(cudaMemsetAsync(op->carg, 0xff, sizeof(long long) * (size_t)op->ncblocks, op->stream));
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_get_courant(RdybOperator op, RdybCourantDiagnostics *d) {
  RDYB_CHECK(op && d, RDYB_ERR_ARG, "NULL argument");
  CU(cudaSetDevice(op->device));
  std::vector<double> v((size_t)op->ncblocks);
  std::vector<long long> a((size_t)op->ncblocks);
  CU(cudaMemcpyAsync(v.data(), op->cmax, sizeof(double) * v.size(), cudaMemcpyDeviceToHost, op->stream));
  CU(cudaMemcpyAsync(a.data(), op->carg, sizeof(long long) * a.size(), cudaMemcpyDeviceToHost, op->stream));
  CU(rdyb_stream_sync(op->stream));
  RDYB_CALL(p2p_check_status(op));
  d->max_courant_num = 0.0; d->global_edge_id = -1; d->global_cell_id = -1; /* ResetOperatorDiagnostics, src/operator.c:757-768 */
  long long arg = -1;
  for (size_t i = 0; i < v.size(); ++i)
    if (v[i] > d->max_courant_num || (v[i] == d->max_courant_num && a[i] >= 0 && (arg < 0 || a[i] < arg))) { d->max_courant_num = v[i]; arg = a[i]; }
  if (arg >= 0 && op->courant_arg_tiled && op->has_tiled) {
    /* tiled kernel: arg = global tile-edge index */
    d->global_edge_id = op->T.edge_id[arg];
    d->global_cell_id = op->L.dev_to_local[op->T.edge_cell[arg]];
  } else if (arg >= 0) {
    /* arg = device cell * ELL width + slot: report the local edge id and the device cell's view-local id; the caller maps
     * them to global ids with RDyEdges/RDyCells.global_ids as the reference does */
    const int32_t cell = (int32_t)(arg / RDYB_ELL_WIDTH), slot = (int32_t)(arg % RDYB_ELL_WIDTH);
    d->global_edge_id = op->L.ell_edge[(size_t)slot * op->L.stride + cell];
    d->global_cell_id = op->L.dev_to_local[cell];
  }
  return RDYB_SUCCESS;
}

/* UpdateOperatorDiagnostics (src/operator.c:867-885): the maximum over all ranks of the communicator.  The reference reduces
 * {value, edge id, cell id} with a custom MPI op; here the value is reduced (ncclAllReduce, max) and the ids are kept on the
 * rank(s) that hold the maximum, -1 elsewhere.  Collective when a communicator exists; the local result otherwise. */
extern "C" int rdyb_operator_get_courant_global(RdybOperator op, RdybCourantDiagnostics *d) {
  RDYB_CALL(rdyb_operator_get_courant(op, d));
  if (!op->comm || op->nranks <= 1) return RDYB_SUCCESS;
  double *scratch = op->partial; /* 1024 doubles of reduction scratch */
  CU(cudaMemcpyAsync(scratch, &d->max_courant_num, sizeof(double), cudaMemcpyHostToDevice, op->stream));
  NC(g_nccl.AllReduce(scratch, scratch + 1, 1, ncclDouble, ncclMax, op->comm, op->stream));
  double global = 0.0;
  CU(cudaMemcpyAsync(&global, scratch + 1, sizeof(double), cudaMemcpyDeviceToHost, op->stream));
  CU(rdyb_stream_sync(op->stream));
  if (global != d->max_courant_num) { d->global_edge_id = -1; d->global_cell_id = -1; }
  d->max_courant_num = global;
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_get_mass(RdybOperator op, double *mass) {
  RDYB_CHECK(op && mass, RDYB_ERR_ARG, "NULL argument");
  CU(cudaSetDevice(op->device));
  RDYB_CALL(adv_resolve(op));
  mass_partial_kernel<<<1024, 256, 0, op->stream>>>(op->state[op->cur], op->area, op->N, op->partial);
  LAUNCHED();
  CU(cudaGetLastError());
  double host[1024];
  CU(cudaMemcpyAsync(host, op->partial, sizeof(host), cudaMemcpyDeviceToHost, op->stream));
  CU(rdyb_stream_sync(op->stream));
  RDYB_CALL(p2p_check_status(op));
  double s = 0.0;
  for (int i = 0; i < 1024; ++i) s += host[i];
  *mass = s;
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_device_pointer(RdybOperator op, int32_t which, void **ptr, int64_t *length) {
  RDYB_CHECK(op && ptr, RDYB_ERR_ARG, "NULL argument");
  RDYB_CHECK(which >= 0 && which < 6, RDYB_ERR_ARG, "unknown array %d", which);
  RDYB_CALL(adv_resolve(op));
  if (which < 3) { *ptr = op->state[op->cur] + (size_t)which * op->stride; if (length) *length = op->NL; }
  else { *ptr = op->F + (size_t)(which - 3) * op->stride; if (length) *length = op->N; }
  return RDYB_SUCCESS;
}

extern "C" int rdyb_operator_last_kernel_ms(RdybOperator op, double *ms_per_launch, int32_t *launches) {
  RDYB_CHECK(op && ms_per_launch, RDYB_ERR_ARG, "NULL argument");
  CU(cudaSetDevice(op->device));
  CU(rdyb_event_sync(op->ev_t1));
  float ms = 0.f;
  CU(cudaEventElapsedTime(&ms, op->ev_t0, op->ev_t1));
  *ms_per_launch = op->last_launches > 0 ? (double)ms / op->last_launches : 0.0;
  if (launches) *launches = op->last_launches;
  return RDYB_SUCCESS;
}

/* which kernel the last RHS / step launch ran (a requested kernel falls back when a mesh does not fit its compiled shapes) */
extern "C" int rdyb_operator_kernel_info(RdybOperator op, char *buf, int32_t len) {
  RDYB_CHECK(op && buf && len > 0, RDYB_ERR_ARG, "NULL argument");
  if (op->kernel_kind >= 1 && op->has_tiled)
    snprintf(buf, (size_t)len, "%s; tiles %d x %d cells, max halo %d, max edges %d; interior cells %d of %d owned, %d local", op->last_kernel, op->T.num_tiles,
             op->T.tile_cells, op->T.max_halo, op->T.max_edges, op->NI, op->N, op->NL);
  else snprintf(buf, (size_t)len, "%s; interior cells %d of %d owned, %d local", op->last_kernel, op->NI, op->N, op->NL);
  return RDYB_SUCCESS;
}

/* pinned host memory for the reference-facing host-Vec path (apply_host) */
extern "C" int rdyb_host_alloc(size_t bytes, void **ptr) {
  RDYB_CHECK(ptr, RDYB_ERR_ARG, "NULL argument");
  CU(cudaMallocHost(ptr, bytes));
  return RDYB_SUCCESS;
}
extern "C" int rdyb_host_free(void *ptr) {
  if (ptr) CU(cudaFreeHost(ptr));
  return RDYB_SUCCESS;
}
#include "swe_forcing.cuh"
