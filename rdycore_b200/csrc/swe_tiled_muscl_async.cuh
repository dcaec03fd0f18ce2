/* swe_tiled_muscl_async.cuh — K3M: the second-order (MUSCL) tile kernel with K3's staging (sm_100a).
 *
 * STATUS (round 2): parity-green on the B200 and 15 - 25 % faster than K2M (10.1 Gcell/s on 128-cell tiles, 11.4 on 192).  Since K4M
 * (swe_tiled_muscl_fixed.cuh: compile-time shape, in-kernel gradients) it is the general fallback: other tile sizes, halos or edge
 * counts beyond K4M's shape, momentum sources.  RDYB_MUSCL_K3M=1 selects it for A/B runs, RDYB_MUSCL_K2M=1 the round-1 kernel.
 *
 * K2M is latency-bound: plain loads in front of two block-wide barriers, hidden only by many small resident blocks
 * (profiles/r1_k2m_sweep.txt).  K3M takes the loads off the compute warps the way K3 does for first order
 * (swe_tiled_async.cuh): persistent CTAs, a producer warp that stages tile n+1 with `cp.async.bulk` (one 1-D bulk copy per
 * contiguous per-tile slice: state, 1/area, the six gradient components, slopes, Manning n, sources, cell-edge slots, edge
 * endpoints / normal / length / the four face offsets) plus 8-byte `cp.async` gathers of the halo cells' ten values, all counted on
 * one mbarrier per stage, while 8 consumer warps compute tile n from shared memory only.  Second order needs no primitive-state
 * pass (velocities follow the reconstruction), so a tile is two phases with ONE consumer barrier:
 *   edges: reconstruct both sides, clamp, Roe once per edge; flux * length is written over the edge's own staged geometry
 *          (sn, cn, length — already in registers), so there is no scratch array and no cross-tile hazard;
 *   cells: signed flux sum in the reference's order, bed slope, friction, sources, F or u + dt*F.
 * Reference functions as in swe_tiled_muscl.cuh (ApplyInteriorFlux2R src/swe/swe_petsc.c:98-213, ReconstructFaceValues
 * src/operator_fluxes_ceed.c:1150-1195).
 */
#ifndef RDYB_SWE_TILED_MUSCL_ASYNC_CUH
#define RDYB_SWE_TILED_MUSCL_ASYNC_CUH
#include "swe_tiled_async.cuh"
#include "swe_tiled_muscl.cuh"

struct MusclAsyncArgs {
  TiledArgs t;
  const double *__restrict__ edge_off; /* [4][plane] */
  size_t plane;                        /* even: every plane starts 16-byte aligned */
  int32_t tile_begin, tile_end;
  int32_t TC, CS, ME;                  /* tile capacity (local index of the first halo cell), staged-cell and edge capacities */
  int32_t stage_bytes;
};

struct MusclStageView {
  double *q[3], *ia, *g[6];                  /* [CS] */
  double *dzdx, *dzdy, *mann, *s0, *s1, *s2; /* [TC] */
  double *esn, *ecn, *elen, *off[4];         /* [ME]; the first three become flux * length after the edge phase */
  uint32_t *elr;                             /* [ME] */
  uint16_t *ce;                              /* [4][TC] */
  int32_t *desc;                             /* [8] */
};

__device__ __forceinline__ MusclStageView muscl_stage_view(char *base, int TC, int CS, int ME) {
  MusclStageView v;
  double *d = reinterpret_cast<double *>(base);
  for (int c = 0; c < 3; ++c) { v.q[c] = d; d += CS; }
  v.ia = d; d += CS;
  for (int c = 0; c < 6; ++c) { v.g[c] = d; d += CS; }
  v.dzdx = d; v.dzdy = v.dzdx + TC; v.mann = v.dzdy + TC; v.s0 = v.mann + TC; v.s1 = v.s0 + TC; v.s2 = v.s1 + TC;
  d = v.s2 + TC;
  v.esn = d; v.ecn = v.esn + ME; v.elen = v.ecn + ME;
  d = v.elen + ME;
  for (int c = 0; c < 4; ++c) { v.off[c] = d; d += ME; }
  v.elr = reinterpret_cast<uint32_t *>(d);
  v.ce = reinterpret_cast<uint16_t *>(v.elr + ME);
  v.desc = reinterpret_cast<int32_t *>(v.ce + 4 * TC);
  return v;
}
/* CS even, TC a multiple of 8, ME a multiple of 4: every array starts on a 16-byte boundary */
__host__ __device__ inline size_t muscl_stage_bytes(int TC, int CS, int ME) {
  return sizeof(double) * (10 * (size_t)CS + 6 * (size_t)TC + 7 * (size_t)ME) + 4 * (size_t)ME + 2 * 4 * (size_t)TC + 32;
}

template <int SRC, bool EULER, int RECON>
__global__ void __launch_bounds__(RDYB_ASYNC_THREADS, 2) swe_tiled_muscl_async_kernel(const MusclAsyncArgs q) {
  extern __shared__ __align__(128) char smem_raw[];
  __shared__ uint64_t full_bar[2], empty_bar[2];
  const TiledArgs &p = q.t;
  const SweArgs &a = p.a;
  const int TC = q.TC, CS = q.CS, ME = q.ME;
  const bool mom = a.s1 != nullptr;
  if (threadIdx.x == 0) {
    mbar_init(&full_bar[0], 33); mbar_init(&full_bar[1], 33); /* 32 lanes' halo gathers + the lane that posts the byte count */
    mbar_init(&empty_bar[0], RDYB_ASYNC_CONSUMERS); mbar_init(&empty_bar[1], RDYB_ASYNC_CONSUMERS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const size_t S = (size_t)a.stride;

  if (warp == RDYB_ASYNC_CONSUMERS / 32) {
    /* ================= producer warp ================= */
    int it = 0;
    for (int tile = q.tile_begin + blockIdx.x; tile < q.tile_end; tile += gridDim.x, ++it) {
      const int s = it & 1;
      mbar_wait(&empty_bar[s], ((it >> 1) & 1) ^ 1);
      const MusclStageView v = muscl_stage_view(smem_raw + (size_t)s * q.stage_bytes, TC, CS, ME);
      const int4 d0 = *reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)tile);
      const int2 d1 = *reinterpret_cast<const int2 *>(p.tile_desc + 8 * (size_t)tile + 4);
      const int cb = d0.x, ncell = d0.y, eo = d0.z, nedge = d0.w, ho = d1.x, nhalo = d1.y;
      if (lane == 0) {
        v.desc[0] = cb; v.desc[1] = ncell; v.desc[2] = eo; v.desc[3] = nedge; v.desc[4] = ho; v.desc[5] = nhalo;
        const uint32_t bc = (uint32_t)((ncell + 1) & ~1) * 8u;  /* per-cell double slices */
        const uint32_t bh = (uint32_t)((ncell + 7) & ~7) * 2u;  /* per-cell uint16 slices */
        const uint32_t be = (uint32_t)((nedge + 1) & ~1) * 8u;  /* per-edge double slices */
        const uint32_t bl = (uint32_t)((nedge + 3) & ~3) * 4u;  /* per-edge uint32 slice */
        const int ncd = 3 + 1 + 6 + 2 + 1 + 1 + (mom ? 2 : 0);
        mbar_expect_tx(&full_bar[s], bc * ncd + bh * a.width + be * 7 + bl);
        bulk_g2s(v.q[0], a.h + cb, bc, &full_bar[s]);
        bulk_g2s(v.q[1], a.hu + cb, bc, &full_bar[s]);
        bulk_g2s(v.q[2], a.hv + cb, bc, &full_bar[s]);
        bulk_g2s(v.ia, a.inv_area + cb, bc, &full_bar[s]);
        for (int c = 0; c < 6; ++c) bulk_g2s(v.g[c], a.grad + (size_t)c * S + cb, bc, &full_bar[s]);
        bulk_g2s(v.dzdx, a.dzdx + cb, bc, &full_bar[s]);
        bulk_g2s(v.dzdy, a.dzdy + cb, bc, &full_bar[s]);
        bulk_g2s(v.mann, a.mann + cb, bc, &full_bar[s]);
        bulk_g2s(v.s0, a.s0 + cb, bc, &full_bar[s]);
        if (mom) { bulk_g2s(v.s1, a.s1 + cb, bc, &full_bar[s]); bulk_g2s(v.s2, a.s2 + cb, bc, &full_bar[s]); }
        for (int k = 0; k < a.width; ++k) bulk_g2s(v.ce + (size_t)k * TC, p.cell_edges + (size_t)k * a.stride + cb, bh, &full_bar[s]);
        bulk_g2s(v.esn, p.edge_sn + eo, be, &full_bar[s]);
        bulk_g2s(v.ecn, p.edge_cn + eo, be, &full_bar[s]);
        bulk_g2s(v.elen, p.edge_len + eo, be, &full_bar[s]);
        for (int c = 0; c < 4; ++c) bulk_g2s(v.off[c], q.edge_off + (size_t)c * q.plane + eo, be, &full_bar[s]);
        bulk_g2s(v.elr, p.edge_lr + eo, bl, &full_bar[s]);
      }
      /* halo gather: 8-byte async copies straight into the staged arrays at local index TC + t */
      for (int t = lane; t < nhalo; t += 32) {
        const int c = p.halo_cells[ho + t];
        cp_async8(v.q[0] + TC + t, a.h + c);
        cp_async8(v.q[1] + TC + t, a.hu + c);
        cp_async8(v.q[2] + TC + t, a.hv + c);
        cp_async8(v.ia + TC + t, a.inv_area + c);
#pragma unroll
        for (int k = 0; k < 6; ++k) cp_async8(v.g[k] + TC + t, a.grad + (size_t)k * S + c);
      }
      cp_async_arrive_noinc(&full_bar[s]);
    }
    return;
  }

  /* ================= consumer warps ================= */
  const double g = RDYB_GRAVITY, tiny = a.tiny_h;
  const int tid = threadIdx.x;
  double cmax = 0.0;
  long long carg = -1;
  int it = 0;
  for (int tile = q.tile_begin + blockIdx.x; tile < q.tile_end; tile += gridDim.x, ++it) {
    const int s = it & 1;
    mbar_wait(&full_bar[s], (it >> 1) & 1);
    const MusclStageView v = muscl_stage_view(smem_raw + (size_t)s * q.stage_bytes, TC, CS, ME);
    const int cb = v.desc[0], ncell = v.desc[1], eo = v.desc[2], nedge = v.desc[3];

    /* ---- edges: one reconstructed flux per tile edge, flux * length written over the edge's staged geometry ---- */
    for (int e = tid; e < nedge; e += RDYB_ASYNC_CONSUMERS) {
      const uint32_t lr = v.elr[e];
      const int l = lr & 0xffffu, r = lr >> 16;
      const double sn = v.esn[e], cn = v.ecn[e], len = v.elen[e];
      RoeFlux f = {0.0, 0.0, 0.0, 0.0};
      if (r != 0xffff) {
        const double dxl = v.off[0][e], dyl = v.off[1][e], dxr = v.off[2][e], dyr = v.off[3][e];
        double fl[3], fr[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const double ql = v.q[c][l], qr = v.q[c][r];
          const double el = v.g[c][l] * dxl + v.g[3 + c][l] * dyl;
          const double er = v.g[c][r] * dxr + v.g[3 + c][r] * dyr;
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
          const double cnum = f.amax * len * fmax(v.ia[l], v.ia[r]) * a.dt;
          if (cnum > cmax) { cmax = cnum; carg = (long long)eo + e; }
        }
      } else {
        /* boundary edge (left cell is a tile cell): first order, non-strict wet test as in ApplyBoundaryFlux */
        const int j = p.edge_bidx[eo + e];
        const double h = v.q[0][l];
        double ul, vl;
        swe_velocities<false>(h, v.q[1][l], v.q[2][l], tiny, a.ha2, ul, vl);
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
          const double cnum = f.amax * len * v.ia[l] * a.dt;
          if (cnum > cmax) { cmax = cnum; carg = (long long)eo + e; }
        }
        if (a.bflux) {
          a.bflux[j] = f.f0; a.bflux[a.nbe + j] = f.f1; a.bflux[2 * a.nbe + j] = f.f2;
          a.bacc[j] += a.dt * f.f0; a.bacc[a.nbe + j] += a.dt * f.f1; a.bacc[2 * a.nbe + j] += a.dt * f.f2;
        }
      }
      v.esn[e] = f.f0 * len; v.ecn[e] = f.f1 * len; v.elen[e] = f.f2 * len; /* this thread's own entries: no one else reads them before the barrier */
    }
    consumer_sync();

    /* ---- cells: flux sum, sources, update ---- */
    for (int t = tid; t < ncell; t += RDYB_ASYNC_CONSUMERS) {
      const int i = cb + t;
      const double h = v.q[0][t], hu = v.q[1][t], hv = v.q[2][t], ia = v.ia[t];
      double F0 = 0.0, F1 = 0.0, F2 = 0.0;
#pragma unroll
      for (int k = 0; k < RDYB_ELL_WIDTH; ++k) {
        if (k >= a.width) break;
        const uint32_t ce = v.ce[k * TC + t];
        if (ce == 0xffffu) continue;
        const int e = ce & 0x7fffu;
        const double sgn = (ce & 0x8000u) ? ia : -ia;
        F0 += v.esn[e] * sgn; F1 += v.ecn[e] * sgn; F2 += v.elen[e] * sgn;
      }
      if (a.fdiv) { a.fdiv[i] = F0; a.fdiv[a.stride + i] = F1; a.fdiv[2 * a.stride + i] = F2; }
      const double bedx = v.dzdx[t] * g * h, bedy = v.dzdy[t] * g * h;
      double tbx = 0.0, tby = 0.0;
      if (h >= tiny) swe_friction<SRC>(h, hu, hv, F1, F2, bedx, bedy, v.mann[t], a.dt, a.xq_thr, tbx, tby);
      const double S1 = mom ? v.s1[t] : 0.0, S2 = mom ? v.s2[t] : 0.0;
      F0 += v.s0[t];
      F1 += -bedx - tbx + S1;
      F2 += -bedy - tby + S2;
      if (a.prim) {
        const bool wet = h >= tiny;
        const double rr = h * swe_rcp(h * h + a.ha2);
        a.prim[i] = h; a.prim[a.stride + i] = wet ? hu * rr : 0.0; a.prim[2 * a.stride + i] = wet ? hv * rr : 0.0;
      }
      if (EULER) { a.o0[i] = h + a.dt * F0; a.o1[i] = hu + a.dt * F1; a.o2[i] = hv + a.dt * F2; }
      else { a.o0[i] = F0; a.o1[i] = F1; a.o2[i] = F2; }
    }
    mbar_arrive(&empty_bar[s]); /* this thread is done reading stage s */
  }

  if (a.cmax) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const double ov = __shfl_down_sync(0xffffffffu, cmax, off);
      const long long oa = __shfl_down_sync(0xffffffffu, carg, off);
      if (ov > cmax || (ov == cmax && oa >= 0 && (carg < 0 || oa < carg))) { cmax = ov; carg = oa; }
    }
    __shared__ double wval[RDYB_ASYNC_CONSUMERS / 32];
    __shared__ long long warg[RDYB_ASYNC_CONSUMERS / 32];
    if (lane == 0) { wval[warp] = cmax; warg[warp] = carg; }
    consumer_sync();
    if (tid == 0) {
      for (int w = 1; w < RDYB_ASYNC_CONSUMERS / 32; ++w)
        if (wval[w] > cmax || (wval[w] == cmax && warg[w] >= 0 && (carg < 0 || warg[w] < carg))) { cmax = wval[w]; carg = warg[w]; }
      const int slot = blockIdx.x; /* one launch covers all tiles: one slot per persistent CTA */
      if (cmax > a.cmax[slot] || (cmax == a.cmax[slot] && carg >= 0 && (a.carg[slot] < 0 || carg < a.carg[slot]))) { a.cmax[slot] = cmax; a.carg[slot] = carg; }
    }
  }
}

/* ---------------- host side ---------------- */
static inline void tiled_muscl_async_shape(const RdybTiledLayout &T, int32_t *CS, int32_t *ME, int32_t *stage_bytes, size_t *smem) {
  *CS = T.tile_cells + ((T.max_halo + 1) & ~1);
  *ME = (T.max_edges + 3) & ~3;
  *stage_bytes = (int32_t)((muscl_stage_bytes(T.tile_cells, *CS, *ME) + 127) & ~(size_t)127);
  *smem = 2 * (size_t)*stage_bytes;
}
/* the two stages of this layout's largest tile fit the 227 KB a block may use, and the face-offset planes are 16-byte aligned */
static inline bool tiled_muscl_async_applicable(const RdybTiledLayout &T, const double *edge_off) {
  int32_t CS, ME, sb;
  size_t smem;
  tiled_muscl_async_shape(T, &CS, &ME, &sb, &smem);
  return edge_off && T.edge_off_plane % 2 == 0 && T.tile_cells % 8 == 0 && smem <= 227 * 1024;
}

template <int SRC, bool EULER, int RECON>
static int tiled_muscl_async_launch_t(const MusclAsyncArgs &q, int ntiles, size_t smem, cudaStream_t stream) {
  static PersistentCache cache = {};
  auto kern = swe_tiled_muscl_async_kernel<SRC, EULER, RECON>;
  int grid = 0;
  RDYB_CALL(persistent_grid(kern, RDYB_ASYNC_THREADS, smem, ntiles, cache, "async MUSCL tile kernel", &grid));
  kern<<<grid, RDYB_ASYNC_THREADS, smem, stream>>>(q);
  return RDYB_SUCCESS;
}

/* all tiles of the rank in one launch; `a.cmax` needs one slot per persistent CTA */
static inline int tiled_muscl_async_launch(const TiledDev &d, const RdybTiledLayout &T, const RdybConfig &cfg, const SweArgs &a, const double *edge_off, bool euler,
                                           cudaStream_t stream) {
  if (T.num_tiles <= 0) return RDYB_SUCCESS;
  MusclAsyncArgs q;
  memset(&q, 0, sizeof(q));
  TiledArgs &p = q.t;
  p.a = a;
  p.tile_desc = d.tile_desc; p.tile_cell_begin = d.tile_cell_begin; p.halo_offsets = d.halo_offsets; p.halo_cells = d.halo_cells; p.edge_offsets = d.edge_offsets;
  p.edge_bidx = d.edge_bidx; p.edge_lr = d.edge_lr; p.edge_sn = d.edge_sn; p.edge_cn = d.edge_cn; p.edge_len = d.edge_len;
  p.cell_edges = d.cell_edges;
  p.max_cells = T.tile_cells;
  q.edge_off = edge_off; q.plane = (size_t)T.edge_off_plane;
  q.tile_begin = 0; q.tile_end = T.num_tiles;
  q.TC = T.tile_cells;
  size_t smem;
  tiled_muscl_async_shape(T, &q.CS, &q.ME, &q.stage_bytes, &smem);
  const bool xq = cfg.source_method == RDYB_SOURCE_IMPLICIT_XQ2018;
  const int nt = T.num_tiles;
#define TM(R) (xq ? (euler ? tiled_muscl_async_launch_t<1, true, R>(q, nt, smem, stream) : tiled_muscl_async_launch_t<1, false, R>(q, nt, smem, stream)) \
                  : (euler ? tiled_muscl_async_launch_t<0, true, R>(q, nt, smem, stream) : tiled_muscl_async_launch_t<0, false, R>(q, nt, smem, stream)))
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
