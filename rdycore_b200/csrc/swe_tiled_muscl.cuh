/* swe_tiled_muscl.cuh — K2M: the second-order (MUSCL) flux + sources + stage update on the tiled layout (sm_100a).
 *
 * Same three phases as K2 (swe_tiled.cuh), one thread block per tile:
 *   1. stage: conserved state (h, hu, hv), 1/area and the six least-squares gradient components of the tile's cells (coalesced) and
 *      of its halo cells (gathered once per tile instead of once per edge slot) into shared memory;
 *   2. edges: one thread per tile edge reconstructs the left and right face states from shared memory
 *      (ReconstructFaceValues, src/operator_fluxes_ceed.c:1150-1195, limiters :1100-1132), clamps the depths at zero and solves the
 *      Roe problem ONCE per edge (ApplyInteriorFlux2R, src/swe/swe_petsc.c:140-200); boundary edges stay first order
 *      (ApplyBoundaryFlux is shared by both orders);
 *   3. cells: signed flux sum in the reference's order, bed slope, friction, external sources, F or u + dt*F.
 * The gradients come from muscl_gradient_kernel (swe_kernels.cuh) after the state halo; on partitioned meshes the ghost cells'
 * gradients arrive through the halo plan before this kernel runs (rhs_pass, swe_device.cu).
 * Against K1<RECON> this halves the Roe solves and replaces 17 scattered 8-byte loads per edge slot by shared-memory reads; the
 * face offsets (midpoint - centroid, left and right) are stored once per tile edge (RdybTiledLayout.edge_off).
 */
#ifndef RDYB_SWE_TILED_MUSCL_CUH
#define RDYB_SWE_TILED_MUSCL_CUH
#include "swe_tiled.cuh"

/* Block shape: measured on the B200 (profiles/r1_k2m_sweep.txt, 10 M triangles, minmod, gradient kernel included): the three
 * phases are latency-bound, so residency decides — 256 threads x 2 blocks/SM (103 registers) 1.79 ms/step, 256 x 3 (80 regs) 1.40,
 * 128 x 4 1.55, 128 x 5 1.33, 128 x 6 1.24, 64 x 12 1.34, 128 x 8 (64 registers, 24 bytes of spill) 1.10 with 128-cell tiles
 * (160-cell tiles 1.25); K1<RECON> 1.45. */
#define RDYB_K2M_THREADS 128
#define RDYB_K2M_MINB 8
#define RDYB_K2M_TILE_CELLS 128 /* the tile size the shape above was tuned for (any tiled layout runs) */

template <int SRC, bool EULER, int RECON, int THREADS = RDYB_K2M_THREADS, int MINB = RDYB_K2M_MINB>
__global__ void __launch_bounds__(THREADS, MINB) swe_tiled_muscl_kernel(const TiledArgs p, const double *__restrict__ edge_off, const size_t edge_plane) {
  extern __shared__ double smem[];
  const SweArgs &a = p.a;
  const int tile = p.tile0 + blockIdx.x;
  const int4 d0 = *reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)tile);
  const int2 d1 = *reinterpret_cast<const int2 *>(p.tile_desc + 8 * (size_t)tile + 4);
  const int cb = d0.x, ncell = d0.y, eo = d0.z, nedge = d0.w, ho = d1.x, nhalo = d1.y;
  const int MS = p.max_staged;
  /* shared memory: 10 arrays over the staged cells (q[3], 1/area, grad[6]), 3 over the tile edges */
  double *sq = smem, *sia = sq + 3 * MS, *sg = sia + MS;
  double *sf0 = sg + 6 * MS, *sf1 = sf0 + p.max_edges, *sf2 = sf1 + p.max_edges;
  const double g = RDYB_GRAVITY, tiny = a.tiny_h;
  const size_t S = (size_t)a.stride;

  /* ---- phase 1: stage tile cells and halo cells ---- */
  for (int t = threadIdx.x; t < ncell + nhalo; t += THREADS) {
    const int c = t < ncell ? cb + t : p.halo_cells[ho + t - ncell];
    const int x = t < ncell ? t : p.max_cells + (t - ncell); /* halo cells sit at local index TC + k */
    sq[x] = a.h[c]; sq[MS + x] = a.hu[c]; sq[2 * MS + x] = a.hv[c];
    sia[x] = a.inv_area[c];
#pragma unroll
    for (int k = 0; k < 6; ++k) sg[k * MS + x] = a.grad[k * S + c];
  }
  __syncthreads();

  /* ---- phase 2: one reconstructed flux per tile edge ---- */
  double cmax = 0.0;
  long long carg = -1;
  for (int e = threadIdx.x; e < nedge; e += THREADS) {
    const size_t ge = (size_t)eo + e;
    const uint32_t lr = p.edge_lr[ge];
    const int l = lr & 0xffffu, r = lr >> 16;
    const double sn = p.edge_sn[ge], cn = p.edge_cn[ge], len = p.edge_len[ge];
    RoeFlux f = {0.0, 0.0, 0.0, 0.0};
    if (r != 0xffff) {
      const double dxl = edge_off[ge], dyl = edge_off[edge_plane + ge], dxr = edge_off[2 * edge_plane + ge], dyr = edge_off[3 * edge_plane + ge];
      double fl[3], fr[3];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const double ql = sq[c * MS + l], qr = sq[c * MS + r];
        const double el = sg[c * MS + l] * dxl + sg[(3 + c) * MS + l] * dyl;
        const double er = sg[c * MS + r] * dxr + sg[(3 + c) * MS + r] * dyr;
        const double dq = qr - ql;
        fl[c] = ql + swe_limit_slope<RECON>(el, 0.5 * dq);
        fr[c] = qr + swe_limit_slope<RECON>(er, -0.5 * dq);
      }
      const double hl = fmax(0.0, fl[0]), hr = fmax(0.0, fr[0]);
      if (!(hl < tiny && hr < tiny)) { /* src/swe/swe_petsc.c:179 */
        double ul, vl, ur, vr;
        swe_velocities<false>(hl, fl[1], fl[2], tiny, a.ha2, ul, vl);
        swe_velocities<false>(hr, fr[1], fr[2], tiny, a.ha2, ur, vr);
        f = swe_roe_flux(hl, ul, vl, swe_sqrt(hl), hr, ur, vr, swe_sqrt(hr), sn, cn);
        const double cnum = f.amax * len * fmax(sia[l], sia[r]) * a.dt;
        if (cnum > cmax) { cmax = cnum; carg = (long long)ge; }
      }
    } else {
      /* boundary edge (left cell is a tile cell): first order, non-strict wet test as in ApplyBoundaryFlux */
      const int j = p.edge_bidx[ge];
      const double h = sq[l];
      double ul, vl;
      swe_velocities<false>(h, sq[MS + l], sq[2 * MS + l], tiny, a.ha2, ul, vl);
      double hr = 0.0, ur = 0.0, vr = 0.0;
      bool ok = true;
      const int type = a.btype[j];
      if (type == RDYB_CONDITION_DIRICHLET) {
        hr = a.bval[j];
        swe_velocities<false>(hr, a.bval[a.nbe + j], a.bval[2 * a.nbe + j], tiny, a.ha2, ur, vr);
      } else if (type == RDYB_CONDITION_REFLECTING) {
        swe_reflecting_bc(h, ul, vl, sn, cn, hr, ur, vr);
      } else {
        ok = swe_critical_outflow_bc(h, ul, vl, sn, cn, hr, ur, vr);
      }
      if (ok && !(h < tiny && hr < tiny)) {
        f = swe_roe_flux(h, ul, vl, swe_sqrt(h), hr, ur, vr, swe_sqrt(hr), sn, cn);
        const double cnum = f.amax * len * sia[l] * a.dt;
        if (cnum > cmax) { cmax = cnum; carg = (long long)ge; }
      }
      if (a.bflux) {
        a.bflux[j] = f.f0; a.bflux[a.nbe + j] = f.f1; a.bflux[2 * a.nbe + j] = f.f2;
        a.bacc[j] += a.dt * f.f0; a.bacc[a.nbe + j] += a.dt * f.f1; a.bacc[2 * a.nbe + j] += a.dt * f.f2;
      }
    }
    sf0[e] = f.f0 * len; sf1[e] = f.f1 * len; sf2[e] = f.f2 * len;
  }
  __syncthreads();

  /* ---- phase 3: per-cell flux sum, sources, update ---- */
  for (int t = threadIdx.x; t < ncell; t += THREADS) {
    const int i = cb + t;
    const double h = sq[t], hu = sq[MS + t], hv = sq[2 * MS + t], ia = sia[t];
    double F0 = 0.0, F1 = 0.0, F2 = 0.0;
#pragma unroll
    for (int k = 0; k < RDYB_ELL_WIDTH; ++k) {
      if (k >= a.width) break;
      const uint32_t ce = p.cell_edges[(size_t)k * a.stride + i];
      if (ce == 0xffffu) continue;
      const int e = ce & 0x7fffu;
      const double s = (ce & 0x8000u) ? ia : -ia; /* +1/A for the right cell, -1/A for the left cell */
      F0 += sf0[e] * s;
      F1 += sf1[e] * s;
      F2 += sf2[e] * s;
    }
    if (a.fdiv) { a.fdiv[i] = F0; a.fdiv[a.stride + i] = F1; a.fdiv[2 * a.stride + i] = F2; }
    const double bedx = a.dzdx[i] * g * h, bedy = a.dzdy[i] * g * h;
    double tbx = 0.0, tby = 0.0;
    if (h >= tiny) swe_friction<SRC>(h, hu, hv, F1, F2, bedx, bedy, a.mann[i], a.dt, a.xq_thr, tbx, tby);
    const double S1 = a.s1 ? a.s1[i] : 0.0, S2 = a.s2 ? a.s2[i] : 0.0;
    F0 += a.s0[i];
    F1 += -bedx - tbx + S1;
    F2 += -bedy - tby + S2;
    if (a.prim) {
      const bool wet = h >= tiny;
      const double rr = h * swe_rcp(h * h + a.ha2);
      a.prim[i] = h; a.prim[a.stride + i] = wet ? hu * rr : 0.0; a.prim[2 * a.stride + i] = wet ? hv * rr : 0.0;
    }
    if (EULER) {
      a.o0[i] = h + a.dt * F0; a.o1[i] = hu + a.dt * F1; a.o2[i] = hv + a.dt * F2;
    } else {
      a.o0[i] = F0; a.o1[i] = F1; a.o2[i] = F2;
    }
  }

  if (a.cmax) { /* uniform across the block */
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const double ov = __shfl_down_sync(0xffffffffu, cmax, off);
      const long long oa = __shfl_down_sync(0xffffffffu, carg, off);
      if (ov > cmax || (ov == cmax && oa >= 0 && (carg < 0 || oa < carg))) { cmax = ov; carg = oa; }
    }
    __shared__ double wval[THREADS / 32];
    __shared__ long long warg[THREADS / 32];
    if ((threadIdx.x & 31) == 0) { wval[threadIdx.x >> 5] = cmax; warg[threadIdx.x >> 5] = carg; }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int w = 1; w < THREADS / 32; ++w)
        if (wval[w] > cmax || (wval[w] == cmax && warg[w] >= 0 && (carg < 0 || warg[w] < carg))) { cmax = wval[w]; carg = warg[w]; }
      if (cmax > a.cmax[tile]) { a.cmax[tile] = cmax; a.carg[tile] = carg; }
    }
  }
}

/* ---------------- host side ---------------- */
static inline size_t tiled_muscl_smem(const RdybTiledLayout &T) {
  return sizeof(double) * (10 * (size_t)(T.tile_cells + T.max_halo) + 3 * (size_t)T.max_edges);
}
/* the tile shapes of this layout fit the 227 KB of shared memory a block may use */
static inline bool tiled_muscl_applicable(const RdybTiledLayout &T, const double *edge_off) { return edge_off && tiled_muscl_smem(T) <= 227 * 1024; }

template <int SRC, bool EULER, int RECON>
static int tiled_muscl_launch_t(const TiledArgs &p, const double *edge_off, size_t plane, int ntiles, size_t smem, cudaStream_t stream) {
  static size_t configured_dev[RDYB_MAX_DEVICES] = {}; /* per instantiation and device */
  int dev_ = 0;
  cudaGetDevice(&dev_);
  size_t &configured = configured_dev[dev_ >= 0 && dev_ < RDYB_MAX_DEVICES ? dev_ : 0];
  if (smem > 48 * 1024 && smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(swe_tiled_muscl_kernel<SRC, EULER, RECON>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return rdyb_fail(RDYB_ERR_CUDA, "tiled MUSCL kernel: %zu bytes of shared memory per tile not available (%s)", smem, cudaGetErrorString(e));
    configured = smem;
  }
  swe_tiled_muscl_kernel<SRC, EULER, RECON><<<ntiles, RDYB_K2M_THREADS, smem, stream>>>(p, edge_off, plane);
  return RDYB_SUCCESS;
}

/* all tiles of the rank in one launch (`a` carries the per-cell arrays, the gradients and, when Courant is on, one slot per tile) */
static inline int tiled_muscl_launch(const TiledDev &d, const RdybTiledLayout &T, const RdybConfig &cfg, const SweArgs &a, const double *edge_off, bool euler,
                                     cudaStream_t stream) {
  if (T.num_tiles <= 0) return RDYB_SUCCESS;
  TiledArgs p;
  memset(&p, 0, sizeof(p));
  p.a = a;
  p.tile_desc = d.tile_desc; p.tile_cell_begin = d.tile_cell_begin; p.halo_offsets = d.halo_offsets; p.halo_cells = d.halo_cells; p.edge_offsets = d.edge_offsets;
  p.edge_bidx = d.edge_bidx; p.edge_lr = d.edge_lr; p.edge_sn = d.edge_sn; p.edge_cn = d.edge_cn; p.edge_len = d.edge_len;
  p.cell_edges = d.cell_edges;
  p.tile0 = 0;
  p.max_cells = T.tile_cells; p.max_staged = T.tile_cells + T.max_halo; p.max_edges = T.max_edges;
  const size_t smem = tiled_muscl_smem(T), plane = (size_t)T.edge_off_plane;
  const bool xq = cfg.source_method == RDYB_SOURCE_IMPLICIT_XQ2018;
  const int nt = T.num_tiles;
#define TM(R) (xq ? (euler ? tiled_muscl_launch_t<1, true, R>(p, edge_off, plane, nt, smem, stream) : tiled_muscl_launch_t<1, false, R>(p, edge_off, plane, nt, smem, stream)) \
                  : (euler ? tiled_muscl_launch_t<0, true, R>(p, edge_off, plane, nt, smem, stream) : tiled_muscl_launch_t<0, false, R>(p, edge_off, plane, nt, smem, stream)))
  int ierr;
  switch (cfg.reconstruction) {
    case RDYB_RECON_MUSCL_MINMOD: ierr = TM(RDYB_RECON_MUSCL_MINMOD); break;
    case RDYB_RECON_MUSCL_NONE: ierr = TM(RDYB_RECON_MUSCL_NONE); break;
    default: ierr = TM(RDYB_RECON_MUSCL_VANLEER); break;
  }
#undef TM
  return ierr;
}
#endif
