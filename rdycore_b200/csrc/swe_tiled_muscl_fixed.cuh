/* swe_tiled_muscl_fixed.cuh — K4M: the second-order (MUSCL) tile kernel with compile-time tile shapes and IN-KERNEL gradients (sm_100a).
 *
 * What the ncu capture of K3M + muscl_gradient_kernel at 10 M cells said (profiles/r2_ncu_full_swe_tiled_muscl_async_v1.txt,
 * r2_ncu_full_muscl_gradient_v1.txt): the step moves 4.0 GB (K3M 246 B/cell, the gradient pass 153 B/cell: it writes 48 B/cell that
 * K3M reads back 1.35 times), 35 % of K3M's stall samples sit at the full-barrier wait (its producer starts the three dependent
 * global loads descriptor -> halo index -> gather only after the stage is free), and 1 036 thread instructions per cell, a third of
 * them address arithmetic on run-time shapes.  K4M changes all three:
 *   - no gradient array, no gradient pass (GM = 2): a stage holds the tile's cells, its halo cells (ring 1) and THEIR outside
 *     neighbours (ring 2, state only).  The least-squares weights of tile and ring-1 cells arrive where the gradients used to (six
 *     planes for triangles), and a first consumer phase turns them into gradients in place with exactly the instruction sequence of
 *     muscl_gradient_kernel (slot order, FMAs): every cell's gradient has the same bits in every tile that stages it, so the two
 *     tiles of a shared edge compute the same flux and the scheme stays conservative to the last bit.
 *     GM = 0 keeps the gradient array (ghost cells of a partitioned mesh get their gradients by halo exchange, and quads have eight
 *     weights per cell, not six): K4M then only brings the two items below.
 *   - the producer prefetches the next tile's descriptor and index lists into registers before it waits for the stage (as K4 does);
 *   - tile shapes are template parameters (TC cells, HM ring-1, H2M ring-2, ME edges): a stage is a struct, every shared-memory
 *     address a constant offset.  Shapes that do not fit run K3M (swe_tiled_muscl_async.cuh), as K4 falls back to K3.
 * The consumer count is the launch's (blockDim.x - 32): 192 threads fill the cell pass of a 192-cell tile exactly.
 * Reference functions as in swe_tiled_muscl.cuh: ComputeLeastSquaresGradients src/operator_fluxes_ceed.c:993-1038,
 * ReconstructFaceValues :1150-1195, ApplyInteriorFlux2R src/swe/swe_petsc.c:98-213.
 */
#ifndef RDYB_SWE_TILED_MUSCL_FIXED_CUH
#define RDYB_SWE_TILED_MUSCL_FIXED_CUH
#include "swe_tiled_muscl_async.cuh"

template <int TC, int HM, int H2M, int ME>
struct MusclFixedStage {
  double q[3][TC + HM + H2M]; /* h, hu, hv of tile | ring 1 | ring 2 */
  double ia[TC + HM];
  double g[6][TC + HM];       /* GM 2: weights {cx0,cx1,cx2,cy0,cy1,cy2} on arrival, gradients {d/dx h,hu,hv, d/dy h,hu,hv} after phase 0 */
  double dzdx[TC], dzdy[TC], mann[TC], s0[TC];
  double esn[ME], ecn[ME], elen[ME], off[4][ME]; /* the first three become flux * length in the edge phase */
  uint32_t elr[ME];
  uint16_t ce[4][TC];
  uint16_t hn[HM][4];         /* staged position of each ring-1 cell's neighbours (device_layout.h, halo_nbr) */
  int32_t desc[8];
};

template <int TC, int HM, int H2M, int ME>
struct MusclFixedSmem {
  MusclFixedStage<TC, HM, H2M, ME> st[2];
  uint64_t full_bar[2], empty_bar[2];
  double wval[8];
  long long warg[8];
};

struct MusclFixedArgs {
  TiledArgs t;
  const double *__restrict__ edge_off; /* [4][plane] */
  size_t plane;
  const int32_t *__restrict__ halo2_cells;
  const uint16_t *__restrict__ halo_nbr; /* [total_halo][4] */
  /* [cells][8] {1/area, cx0, cx1, cx2, cy0, cy1, cy2, 0}: what a ring-1 cell needs besides its state, packed into one 64-byte record.
   * Gathered as seven 8-byte copies they touch 2 sectors of 32 bytes instead of the 7 that the separate planes cost (cp.async.ca: the
   * copies of one lane after the first hit L1); NULL: read the planes */
  const double *__restrict__ halo_rec;
  int32_t tile_begin, tile_end;
};

__device__ __forceinline__ void muscl_consumer_sync(int nc) { asm volatile("bar.sync 1, %0;" ::"r"(nc) : "memory"); }

template <int TC, int HM, int H2M, int ME, int SRC, bool EULER, int RECON, int GM>
__global__ void __launch_bounds__(288, 2) swe_tiled_muscl_fixed_kernel(const MusclFixedArgs q) {
  extern __shared__ __align__(128) char smem_raw[];
  using Smem = MusclFixedSmem<TC, HM, H2M, ME>;
  Smem &S = *reinterpret_cast<Smem *>(smem_raw);
  const TiledArgs &p = q.t;
  const SweArgs &a = p.a;
  const int NC = (int)blockDim.x - 32; /* consumer threads */
  constexpr int HK = (HM + 31) / 32, H2K = (H2M + 31) / 32;
  if (threadIdx.x == 0) {
    mbar_init(&S.full_bar[0], 33); mbar_init(&S.full_bar[1], 33); /* 32 lanes' gathers + the lane that posts the byte count */
    mbar_init(&S.empty_bar[0], NC); mbar_init(&S.empty_bar[1], NC);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const size_t Sd = (size_t)a.stride;
  const int G = gridDim.x;

  if (warp == NC / 32) {
    /* ================= producer warp ================= */
    int tile = q.tile_begin + blockIdx.x;
    int4 d0 = make_int4(0, 0, 0, 0), d1 = d0, n0 = d0, n1 = d0;
    if (tile < q.tile_end) {
      d0 = __ldg(reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)tile));
      d1 = __ldg(reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)tile + 4));
    }
    if (tile + G < q.tile_end) {
      n0 = __ldg(reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)(tile + G)));
      n1 = __ldg(reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)(tile + G) + 4));
    }
    int hidx[HK], h2idx[H2K];
#pragma unroll
    for (int k = 0; k < HK; ++k) { const int t = lane + 32 * k; hidx[k] = t < d1.y ? __ldg(p.halo_cells + d1.x + t) : 0; }
#pragma unroll
    for (int k = 0; k < H2K; ++k) { const int t = lane + 32 * k; h2idx[k] = (GM == 2 && t < d1.w) ? __ldg(q.halo2_cells + d1.z + t) : 0; }
    for (int it = 0; tile < q.tile_end; tile += G, ++it) {
      const int s = it & 1;
      /* prefetch: descriptor two tiles ahead, index lists one tile ahead */
      int4 nn0 = make_int4(0, 0, 0, 0), nn1 = nn0;
      if (tile + 2 * G < q.tile_end) {
        nn0 = __ldg(reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)(tile + 2 * G)));
        nn1 = __ldg(reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)(tile + 2 * G) + 4));
      }
      int nh[HK], nh2[H2K];
#pragma unroll
      for (int k = 0; k < HK; ++k) { const int t = lane + 32 * k; nh[k] = t < n1.y ? __ldg(p.halo_cells + n1.x + t) : 0; }
#pragma unroll
      for (int k = 0; k < H2K; ++k) { const int t = lane + 32 * k; nh2[k] = (GM == 2 && t < n1.w) ? __ldg(q.halo2_cells + n1.z + t) : 0; }
      mbar_wait(&S.empty_bar[s], ((it >> 1) & 1) ^ 1);
      auto &st = S.st[s];
      uint64_t *bar = &S.full_bar[s];
      const int cb = d0.x, ncell = d0.y, eo = d0.z, nedge = d0.w, ho = d1.x, nhalo = d1.y, nhalo2 = GM == 2 ? d1.w : 0;
      if (lane == 0) {
        st.desc[0] = cb; st.desc[1] = ncell; st.desc[2] = eo; st.desc[3] = nedge; st.desc[5] = nhalo;
        const uint32_t bc = (uint32_t)((ncell + 1) & ~1) * 8u; /* per-cell double slices */
        const uint32_t bh = (uint32_t)((ncell + 7) & ~7) * 2u; /* per-cell uint16 slices */
        const uint32_t be = (uint32_t)((nedge + 1) & ~1) * 8u; /* per-edge double slices */
        const uint32_t bl = (uint32_t)((nedge + 3) & ~3) * 4u; /* per-edge uint32 slice */
        const uint32_t bn = GM == 2 ? (uint32_t)((nhalo + 1) & ~1) * 8u : 0u; /* ring-1 neighbour positions, 8 bytes per halo cell */
        mbar_expect_tx(bar, bc * 14 + bh * a.width + be * 7 + bl + bn);
        bulk_g2s(st.q[0], a.h + cb, bc, bar);
        bulk_g2s(st.q[1], a.hu + cb, bc, bar);
        bulk_g2s(st.q[2], a.hv + cb, bc, bar);
        bulk_g2s(st.ia, a.inv_area + cb, bc, bar);
        if (GM == 2) {
#pragma unroll
          for (int k = 0; k < 3; ++k) { bulk_g2s(st.g[k], a.gcx + (size_t)k * Sd + cb, bc, bar); bulk_g2s(st.g[3 + k], a.gcy + (size_t)k * Sd + cb, bc, bar); }
          if (bn) bulk_g2s(st.hn, q.halo_nbr + 4 * (size_t)ho, bn, bar);
        } else {
#pragma unroll
          for (int c = 0; c < 6; ++c) bulk_g2s(st.g[c], a.grad + (size_t)c * Sd + cb, bc, bar);
        }
        bulk_g2s(st.dzdx, a.dzdx + cb, bc, bar);
        bulk_g2s(st.dzdy, a.dzdy + cb, bc, bar);
        bulk_g2s(st.mann, a.mann + cb, bc, bar);
        bulk_g2s(st.s0, a.s0 + cb, bc, bar);
        for (int k = 0; k < a.width; ++k) bulk_g2s(st.ce[k], p.cell_edges + (size_t)k * Sd + cb, bh, bar);
        bulk_g2s(st.esn, p.edge_sn + eo, be, bar);
        bulk_g2s(st.ecn, p.edge_cn + eo, be, bar);
        bulk_g2s(st.elen, p.edge_len + eo, be, bar);
#pragma unroll
        for (int c = 0; c < 4; ++c) bulk_g2s(st.off[c], q.edge_off + (size_t)c * q.plane + eo, be, bar);
        bulk_g2s(st.elr, p.edge_lr + eo, bl, bar);
      }
      /* ring 1: state, 1/area and the six weights (GM 2) or gradients (GM 0), 8-byte async copies to staged index TC + t */
#pragma unroll
      for (int k = 0; k < HK; ++k) {
        const int t = lane + 32 * k;
        if (t < nhalo) {
          const int c = hidx[k];
          cp_async8(&st.q[0][TC + t], a.h + c);
          cp_async8(&st.q[1][TC + t], a.hu + c);
          cp_async8(&st.q[2][TC + t], a.hv + c);
          if (GM == 2 && q.halo_rec) {
            const double *rec = q.halo_rec + 8 * (size_t)c;
            cp_async8(&st.ia[TC + t], rec);
#pragma unroll
            for (int j = 0; j < 6; ++j) cp_async8(&st.g[j][TC + t], rec + 1 + j);
          } else if (GM == 2) {
            cp_async8(&st.ia[TC + t], a.inv_area + c);
#pragma unroll
            for (int j = 0; j < 3; ++j) { cp_async8(&st.g[j][TC + t], a.gcx + (size_t)j * Sd + c); cp_async8(&st.g[3 + j][TC + t], a.gcy + (size_t)j * Sd + c); }
          } else {
            cp_async8(&st.ia[TC + t], a.inv_area + c);
#pragma unroll
            for (int j = 0; j < 6; ++j) cp_async8(&st.g[j][TC + t], a.grad + (size_t)j * Sd + c);
          }
        }
      }
      /* ring 2: state only, at staged index TC + HM + t */
      if (GM == 2) {
#pragma unroll
        for (int k = 0; k < H2K; ++k) {
          const int t = lane + 32 * k;
          if (t < nhalo2) {
            const int c = h2idx[k];
            cp_async8(&st.q[0][TC + HM + t], a.h + c);
            cp_async8(&st.q[1][TC + HM + t], a.hu + c);
            cp_async8(&st.q[2][TC + HM + t], a.hv + c);
          }
        }
      }
      cp_async_arrive_noinc(bar);
      d0 = n0; d1 = n1; n0 = nn0; n1 = nn1;
#pragma unroll
      for (int k = 0; k < HK; ++k) hidx[k] = nh[k];
#pragma unroll
      for (int k = 0; k < H2K; ++k) h2idx[k] = nh2[k];
    }
    return;
  }

  /* ================= consumer warps ================= */
  const double g = RDYB_GRAVITY, tiny = a.tiny_h, ha2 = a.ha2, dt = a.dt;
  const int tid = threadIdx.x;
  double cmax = 0.0;
  long long carg = -1;
  int it = 0;
  for (int tile = q.tile_begin + blockIdx.x; tile < q.tile_end; tile += G, ++it) {
    const int s = it & 1;
    mbar_wait(&S.full_bar[s], (it >> 1) & 1);
    auto &st = S.st[s];
    const int cb = st.desc[0], ncell = st.desc[1], eo = st.desc[2], nedge = st.desc[3], nhalo = st.desc[5];

    if (GM == 2) {
      /* ---- phase 0: weights -> gradients, in place.  The sequence of muscl_gradient_kernel (swe_kernels.cuh): slots in ELL order,
       * boundary and empty slots skipped, one FMA per component and direction.  A thread reads the six weights of ITS cell and then
       * overwrites them; other threads only read states in this phase. ---- */
      for (int t = tid; t < ncell; t += NC) {
        const double q0 = st.q[0][t], q1 = st.q[1][t], q2 = st.q[2][t];
        double gx0 = 0.0, gx1 = 0.0, gx2 = 0.0, gy0 = 0.0, gy1 = 0.0, gy2 = 0.0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const uint32_t ce = st.ce[k][t];
          if (ce == 0xffffu) continue;
          const uint32_t lr = st.elr[ce & 0x7fffu];
          if ((lr >> 16) == 0xffffu) continue; /* boundary edge */
          const int n = (ce & 0x8000u) ? (int)(lr & 0xffffu) : (int)(lr >> 16);
          const double cx = st.g[k][t], cy = st.g[3 + k][t];
          const double d0 = st.q[0][n] - q0, d1 = st.q[1][n] - q1, d2 = st.q[2][n] - q2;
          gx0 = fma(cx, d0, gx0); gy0 = fma(cy, d0, gy0); gx1 = fma(cx, d1, gx1); gy1 = fma(cy, d1, gy1); gx2 = fma(cx, d2, gx2); gy2 = fma(cy, d2, gy2);
        }
        st.g[0][t] = gx0; st.g[1][t] = gx1; st.g[2][t] = gx2; st.g[3][t] = gy0; st.g[4][t] = gy1; st.g[5][t] = gy2;
      }
      for (int j = tid; j < nhalo; j += NC) {
        const int x = TC + j;
        const double q0 = st.q[0][x], q1 = st.q[1][x], q2 = st.q[2][x];
        const uint2 hv = *reinterpret_cast<const uint2 *>(st.hn[j]);
        double gx0 = 0.0, gx1 = 0.0, gx2 = 0.0, gy0 = 0.0, gy1 = 0.0, gy2 = 0.0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          const uint32_t v = k == 0 ? (hv.x & 0xffffu) : k == 1 ? (hv.x >> 16) : (hv.y & 0xffffu);
          if (v == 0xffffu) continue;
          const int n = (v & 0x8000u) ? TC + HM + (int)(v & 0x7fffu) : (int)v;
          const double cx = st.g[k][x], cy = st.g[3 + k][x];
          const double d0 = st.q[0][n] - q0, d1 = st.q[1][n] - q1, d2 = st.q[2][n] - q2;
          gx0 = fma(cx, d0, gx0); gy0 = fma(cy, d0, gy0); gx1 = fma(cx, d1, gx1); gy1 = fma(cy, d1, gy1); gx2 = fma(cx, d2, gx2); gy2 = fma(cy, d2, gy2);
        }
        st.g[0][x] = gx0; st.g[1][x] = gx1; st.g[2][x] = gx2; st.g[3][x] = gy0; st.g[4][x] = gy1; st.g[5][x] = gy2;
      }
      muscl_consumer_sync(NC);
    }

    /* ---- edges: one reconstructed flux per tile edge, flux * length written over the edge's staged geometry ---- */
    for (int e = tid; e < nedge; e += NC) {
      const uint32_t lr = st.elr[e];
      const int l = lr & 0xffffu, r = lr >> 16;
      const double sn = st.esn[e], cn = st.ecn[e], len = st.elen[e];
      RoeFlux f = {0.0, 0.0, 0.0, 0.0};
      if (r != 0xffff) {
        const double dxl = st.off[0][e], dyl = st.off[1][e], dxr = st.off[2][e], dyr = st.off[3][e];
        double fl[3], fr[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const double ql = st.q[c][l], qr = st.q[c][r];
          const double el = st.g[c][l] * dxl + st.g[3 + c][l] * dyl;
          const double er = st.g[c][r] * dxr + st.g[3 + c][r] * dyr;
          const double dq = qr - ql;
          fl[c] = ql + swe_limit_slope<RECON>(el, 0.5 * dq);
          fr[c] = qr + swe_limit_slope<RECON>(er, -0.5 * dq);
        }
        const double hl = fmax(0.0, fl[0]), hr = fmax(0.0, fr[0]);
        if (!(hl < tiny && hr < tiny)) { /* src/swe/swe_petsc.c:179 */
          double ul, vl, sql, ur, vr, sqr;
          swe_edge_side(hl, fl[1], fl[2], tiny, ha2, ul, vl, sql);
          swe_edge_side(hr, fr[1], fr[2], tiny, ha2, ur, vr, sqr);
          f = swe_roe_flux(hl, ul, vl, sql, hr, ur, vr, sqr, sn, cn);
          const double cnum = f.amax * len * fmax(st.ia[l], st.ia[r]) * dt;
          if (cnum > cmax) { cmax = cnum; carg = (long long)eo + e; }
        }
      } else {
        /* boundary edge (left cell is a tile cell): first order, non-strict wet test as in ApplyBoundaryFlux */
        const int j = p.edge_bidx[eo + e];
        const double h = st.q[0][l];
        double ul, vl;
        swe_velocities<false>(h, st.q[1][l], st.q[2][l], tiny, ha2, ul, vl);
        double hr = 0.0, ur = 0.0, vr = 0.0;
        bool ok = true;
        const int type = a.btype[j];
        if (type == RDYB_CONDITION_DIRICHLET) {
          hr = a.bval[j];
          swe_velocities<false>(hr, a.bval[a.nbe + j], a.bval[2 * a.nbe + j], tiny, ha2, ur, vr);
        } else if (type == RDYB_CONDITION_REFLECTING) {
          swe_reflecting_bc(h, ul, vl, sn, cn, hr, ur, vr);
        } else {
          ok = swe_critical_outflow_bc(h, ul, vl, sn, cn, hr, ur, vr);
        }
        if (ok && !(h < tiny && hr < tiny)) {
          f = swe_roe_flux(h, ul, vl, swe_sqrt(h), hr, ur, vr, swe_sqrt(hr), sn, cn);
          const double cnum = f.amax * len * st.ia[l] * dt;
          if (cnum > cmax) { cmax = cnum; carg = (long long)eo + e; }
        }
        if (a.bflux) {
          a.bflux[j] = f.f0; a.bflux[a.nbe + j] = f.f1; a.bflux[2 * a.nbe + j] = f.f2;
          a.bacc[j] += dt * f.f0; a.bacc[a.nbe + j] += dt * f.f1; a.bacc[2 * a.nbe + j] += dt * f.f2;
        }
      }
      st.esn[e] = f.f0 * len; st.ecn[e] = f.f1 * len; st.elen[e] = f.f2 * len; /* this thread's own entries: no one else reads them before the barrier */
    }
    muscl_consumer_sync(NC);

    /* ---- cells: flux sum, sources, update ---- */
    for (int t = tid; t < ncell; t += NC) {
      const int i = cb + t;
      const double h = st.q[0][t], hu = st.q[1][t], hv = st.q[2][t], ia = st.ia[t];
      double F0 = 0.0, F1 = 0.0, F2 = 0.0;
#pragma unroll
      for (int k = 0; k < RDYB_ELL_WIDTH; ++k) {
        if (k >= a.width) break;
        const uint32_t ce = st.ce[k][t];
        if (ce == 0xffffu) continue;
        const int e = ce & 0x7fffu;
        const double sgn = (ce & 0x8000u) ? ia : -ia;
        F0 += st.esn[e] * sgn; F1 += st.ecn[e] * sgn; F2 += st.elen[e] * sgn;
      }
      if (a.fdiv) { a.fdiv[i] = F0; a.fdiv[a.stride + i] = F1; a.fdiv[2 * a.stride + i] = F2; }
      const double bedx = st.dzdx[t] * g * h, bedy = st.dzdy[t] * g * h;
      double tbx = 0.0, tby = 0.0;
      if (h >= tiny) swe_friction<SRC>(h, hu, hv, F1, F2, bedx, bedy, st.mann[t], dt, a.xq_thr, tbx, tby);
      F0 += st.s0[t];
      F1 += -bedx - tbx;
      F2 += -bedy - tby;
      if (a.prim) {
        const bool wet = h >= tiny;
        const double rr = h * swe_rcp(h * h + ha2);
        a.prim[i] = h; a.prim[a.stride + i] = wet ? hu * rr : 0.0; a.prim[2 * a.stride + i] = wet ? hv * rr : 0.0;
      }
      if (EULER) { a.o0[i] = h + dt * F0; a.o1[i] = hu + dt * F1; a.o2[i] = hv + dt * F2; }
      else { a.o0[i] = F0; a.o1[i] = F1; a.o2[i] = F2; }
    }
    mbar_arrive(&S.empty_bar[s]); /* this thread is done reading stage s */
  }

  if (a.cmax) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const double ov = __shfl_down_sync(0xffffffffu, cmax, off);
      const long long oa = __shfl_down_sync(0xffffffffu, carg, off);
      if (ov > cmax || (ov == cmax && oa >= 0 && (carg < 0 || oa < carg))) { cmax = ov; carg = oa; }
    }
    if (lane == 0) { S.wval[warp] = cmax; S.warg[warp] = carg; }
    muscl_consumer_sync(NC);
    if (tid == 0) {
      for (int w = 1; w < NC / 32; ++w)
        if (S.wval[w] > cmax || (S.wval[w] == cmax && S.warg[w] >= 0 && (carg < 0 || S.warg[w] < carg))) { cmax = S.wval[w]; carg = S.warg[w]; }
      const int slot = blockIdx.x; /* one launch covers all tiles: one slot per persistent CTA */
      if (cmax > a.cmax[slot] || (cmax == a.cmax[slot] && carg >= 0 && (a.carg[slot] < 0 || carg < a.carg[slot]))) { a.cmax[slot] = cmax; a.carg[slot] = carg; }
    }
  }
}

/* ---------------- host side ---------------- */
#define RDYB_K4M_TC 192
#define RDYB_K4M_HM 80
#define RDYB_K4M_H2M 96
#define RDYB_K4M_ME 352

/* gradient mode this layout and these arguments can run: 2 = in-kernel gradients (triangles, every ring-1 cell owned), 0 = gradient
 * array, -1 = not K4M's shape (K3M runs it) */
static inline int tiled_muscl_fixed_mode(const RdybTiledLayout &T, const double *edge_off, const SweArgs &a, bool two_ring_on_device) {
  if (!edge_off || T.edge_off_plane % 2 != 0 || T.tile_cells != RDYB_K4M_TC || T.max_halo > RDYB_K4M_HM || T.max_edges > RDYB_K4M_ME) return -1;
  if (a.s1 || a.s2) return -1; /* momentum sources: K3M */
  if (getenv("RDYB_MUSCL_K3M")) return -1;
  const bool two = T.two_ring && two_ring_on_device && a.width == 3 && T.max_halo2 <= RDYB_K4M_H2M && !getenv("RDYB_K4M_GRAD_ARRAY");
  return two ? 2 : 0;
}
/* consumer threads per CTA: 192 fill the cell pass of a 192-cell tile exactly (and the ~325-edge pass to 85 %); RDYB_K4M_CONSUMERS=256 for A/B */
static inline int tiled_muscl_fixed_threads() {
  const char *env = getenv("RDYB_K4M_CONSUMERS");
  const int nc = env ? atoi(env) : 192;
  return (nc == 128 || nc == 160 || nc == 192 || nc == 224 || nc == 256 ? nc : 192) + 32;
}

template <int SRC, bool EULER, int RECON, int GM>
static int tiled_muscl_fixed_launch_t(const MusclFixedArgs &q, int ntiles, cudaStream_t stream) {
  static PersistentCache cache = {};
  auto kern = swe_tiled_muscl_fixed_kernel<RDYB_K4M_TC, RDYB_K4M_HM, RDYB_K4M_H2M, RDYB_K4M_ME, SRC, EULER, RECON, GM>;
  const size_t smem = sizeof(MusclFixedSmem<RDYB_K4M_TC, RDYB_K4M_HM, RDYB_K4M_H2M, RDYB_K4M_ME>);
  const int threads = tiled_muscl_fixed_threads();
  int grid = 0;
  RDYB_CALL(persistent_grid(kern, threads, smem, ntiles, cache, "fixed-shape MUSCL tile kernel", &grid));
  kern<<<grid, threads, smem, stream>>>(q);
  return RDYB_SUCCESS;
}

/* all tiles of the rank in one launch; `a.cmax` needs one slot per persistent CTA.  mode from tiled_muscl_fixed_mode (0 or 2) */
static inline int tiled_muscl_fixed_launch(const TiledDev &d, const RdybTiledLayout &T, const RdybConfig &cfg, const SweArgs &a, const double *edge_off,
                                           const int32_t *halo2_cells, const uint16_t *halo_nbr, const double *halo_rec, int mode, bool euler,
                                           cudaStream_t stream) {
  if (T.num_tiles <= 0) return RDYB_SUCCESS;
  MusclFixedArgs q;
  memset(&q, 0, sizeof(q));
  TiledArgs &p = q.t;
  p.a = a;
  p.tile_desc = d.tile_desc; p.tile_cell_begin = d.tile_cell_begin; p.halo_offsets = d.halo_offsets; p.halo_cells = d.halo_cells; p.edge_offsets = d.edge_offsets;
  p.edge_bidx = d.edge_bidx; p.edge_lr = d.edge_lr; p.edge_sn = d.edge_sn; p.edge_cn = d.edge_cn; p.edge_len = d.edge_len;
  p.cell_edges = d.cell_edges;
  p.max_cells = T.tile_cells;
  q.edge_off = edge_off; q.plane = (size_t)T.edge_off_plane;
  q.halo2_cells = halo2_cells; q.halo_nbr = halo_nbr; q.halo_rec = getenv("RDYB_K4M_NO_HALO_REC") ? nullptr : halo_rec;
  q.tile_begin = 0; q.tile_end = T.num_tiles;
  const bool xq = cfg.source_method == RDYB_SOURCE_IMPLICIT_XQ2018;
  const int nt = T.num_tiles;
#define TMG(R, GMv) (xq ? (euler ? tiled_muscl_fixed_launch_t<1, true, R, GMv>(q, nt, stream) : tiled_muscl_fixed_launch_t<1, false, R, GMv>(q, nt, stream)) \
                        : (euler ? tiled_muscl_fixed_launch_t<0, true, R, GMv>(q, nt, stream) : tiled_muscl_fixed_launch_t<0, false, R, GMv>(q, nt, stream)))
#define TM(R) (mode == 2 ? TMG(R, 2) : TMG(R, 0))
  int ierr;
  switch (cfg.reconstruction) {
    case RDYB_RECON_MUSCL_MINMOD: ierr = TM(RDYB_RECON_MUSCL_MINMOD); break;
    case RDYB_RECON_MUSCL_NONE: ierr = TM(RDYB_RECON_MUSCL_NONE); break;
    default: ierr = TM(RDYB_RECON_MUSCL_VANLEER); break;
  }
#undef TM
#undef TMG
  return ierr;
}
#endif
