/* swe_tiled_split.cuh — K6: K4's tile pipeline without a producer warp and with the per-tile work split ACROSS threads.
 *
 * What the ncu source page of K4 (profiles/r1_ncu_full_swe_tiled_fixed.txt) showed: 17 % of all warp samples sit at the two
 * CTA barriers of a tile.  K4 gives its 224 consumer threads ~370 edges (two passes, the second 2/3 empty), then 224 cells,
 * then ~285 primitive states (two passes, the second 1/4 full): a warp waits for the longest pass of every phase, and the
 * producer warp holds 1/8 of the CTA's registers while it sleeps.
 *
 * K6 keeps K4's device layout, staging (cp.async.bulk + mbarrier) and arithmetic, and changes the schedule:
 *   - 384 threads, 2 CTAs per SM (24 compute warps per SM instead of 21), three stages in shared memory;
 *   - region A of tile k: ONE edge per thread (Roe flux, boundary branch, Courant candidate);
 *   - region B of tile k: threads [0,TC) update one cell of tile k each (flux sum, sources, stage update) WHILE threads
 *     [TC,384) compute the primitive states of tile k+1 — the two are independent, so nobody idles;
 *   - no producer warp: at the top of iteration k every warp issues ONE of the bulk copies of tile k+2 and the first 96 threads
 *     gather one halo cell each.  Nothing is kept in registers across iterations, because the tile's descriptor and halo index
 *     list arrive as one 352-byte RECORD (tile_rec, built once per layout) by a bulk copy issued one iteration earlier.
 * Two CTA barriers per tile remain, but every thread has the same amount of work between them.
 *
 * Replaces (fused) the same reference functions as K4: ApplyInteriorFlux src/swe/swe_petsc.c:215-316, ApplyInteriorFluxHR
 * :1000-1161, ApplyBoundaryFlux :506-630, ApplySourceSemiImplicit :704-804, ApplySourceImplicitXQ2018 :816-932, TSEULER.
 */
#ifndef RDYB_SWE_TILED_SPLIT_CUH
#define RDYB_SWE_TILED_SPLIT_CUH
#include "swe_tiled_fixed.cuh"

#define RDYB_SPLIT_REC_HALO 80                        /* halo capacity of a record */
#define RDYB_SPLIT_REC_INTS (8 + RDYB_SPLIT_REC_HALO) /* {tile_desc[8], halo cell indices}: 352 bytes, a multiple of 16 */

template <int TC, int HM, int ME, bool HR>
struct SplitSmem {
  FixedStage<TC, HM, ME, HR> st[3];
  FixedScratch<TC, HM, ME, HR, false> sc;
  alignas(16) int32_t rec[4][RDYB_SPLIT_REC_INTS];
  uint64_t full_bar[3], rec_bar[4];
  double wval[16];
  int wargs[16];
};

__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

/* ALL threads: start the copies of this CTA's j-th tile into stage j % 3 (its record has been prefetched).  The work is spread
 * over the CTA — lane 0 of warp w issues bulk copy number w (a first version let one warp issue all 14 copies and the halo gather:
 * ~265 dependent instructions, 1/3 of that warp's time per tile, and 30 % of all warp samples ended up at the barrier waiting for
 * it, profiles/r1_ncu_k6_v1.txt).  Thread 0 alone posts the byte count: its arrive.expect_tx is one of the counted arrivals, so the
 * phase cannot complete before the count is known even if other copies land first. */
#define RDYB_SPLIT_GATHER_THREADS 96
template <int TC, int HM, int ME, bool HR>
__device__ __forceinline__ void split_issue_tile(SplitSmem<TC, HM, ME, HR> &S, const TiledArgs &p, int j, int tid) {
  static_assert(HM <= RDYB_SPLIT_GATHER_THREADS, "one gather thread per halo cell");
  const SweArgs &a = p.a;
  mbar_wait(&S.rec_bar[j & 3], (j >> 2) & 1);
  const int32_t *rec = S.rec[j & 3];
  auto &st = S.st[j % 3];
  uint64_t *bar = &S.full_bar[j % 3];
  if ((tid & 31) == 0) {
    const int cb = rec[0], ncell = rec[1], eo = rec[2], nedge = rec[3];
    const uint32_t bc = (uint32_t)((ncell + 1) & ~1) * 8u; /* per-cell double slices */
    const uint32_t bh = (uint32_t)((ncell + 7) & ~7) * 2u; /* per-cell uint16 slices */
    const uint32_t be = (uint32_t)((nedge + 1) & ~1) * 8u; /* per-edge double slices */
    const uint32_t bl = (uint32_t)((nedge + 3) & ~3) * 4u; /* per-edge uint32 slice */
    const int w = tid >> 5, width = a.width;
    switch (w) {
      case 0: {
        st.desc[0] = cb; st.desc[1] = ncell; st.desc[2] = eo; st.desc[3] = nedge; st.desc[5] = rec[5];
        constexpr int ncd = 3 + 1 + (HR ? 1 : 2) + 1 + 1;
        mbar_expect_tx(bar, bc * ncd + bh * width + be * 3 + bl);
        bulk_g2s(st.h, a.h + cb, bc, bar);
      } break;
      case 1: bulk_g2s(st.hu, a.hu + cb, bc, bar); break;
      case 2: bulk_g2s(st.hv, a.hv + cb, bc, bar); break;
      case 3: bulk_g2s(st.ia, a.inv_area + cb, bc, bar); break;
      case 4:
        if (HR) bulk_g2s(st.z, a.zc + cb, bc, bar); else bulk_g2s(st.dzdx, a.dzdx + cb, bc, bar);
        if (width > 0) bulk_g2s(st.ce[0], p.cell_edges + cb, bh, bar);
        break;
      case 5:
        if (!HR) bulk_g2s(st.dzdy, a.dzdy + cb, bc, bar);
        if (width > 1) bulk_g2s(st.ce[1], p.cell_edges + (size_t)a.stride + cb, bh, bar);
        break;
      case 6:
        bulk_g2s(st.mann, a.mann + cb, bc, bar);
        if (width > 2) bulk_g2s(st.ce[2], p.cell_edges + 2 * (size_t)a.stride + cb, bh, bar);
        break;
      case 7:
        bulk_g2s(st.s0, a.s0 + cb, bc, bar);
        if (width > 3) bulk_g2s(st.ce[3], p.cell_edges + 3 * (size_t)a.stride + cb, bh, bar);
        break;
      case 8: bulk_g2s(st.esn, p.edge_sn + eo, be, bar); break;
      case 9: bulk_g2s(st.ecn, p.edge_cn + eo, be, bar); break;
      case 10: bulk_g2s(st.elen, p.edge_len + eo, be, bar); break;
      default: bulk_g2s(st.elr, p.edge_lr + eo, bl, bar); break;
    }
  }
  /* halo gather: 8-byte async copies straight into the staged arrays at local index TC + t, one halo cell per thread */
  if (tid < RDYB_SPLIT_GATHER_THREADS) {
    if (tid < rec[5]) {
      const int c = rec[8 + tid];
      cp_async8(st.h + TC + tid, a.h + c);
      cp_async8(st.hu + TC + tid, a.hu + c);
      cp_async8(st.hv + TC + tid, a.hv + c);
      cp_async8(st.ia + TC + tid, a.inv_area + c);
      if constexpr (HR) cp_async8(st.z + TC + tid, a.zc + c);
    }
    cp_async_arrive_noinc(bar);
  }
}

/* primitive state (u, v, sqrt h) of staged cell x: ComputeRiemannVelocities, src/swe/swe_petsc.c:57-73 */
template <bool HR, class Stage, class Scratch>
__device__ __forceinline__ void split_prim(const Stage &st, Scratch &sc, int x, double tiny, double ha2) {
  const double h = st.h[x];
  double u, w;
  swe_velocities<HR>(h, st.hu[x], st.hv[x], tiny, ha2, u, w);
  sc.u[x] = u; sc.v[x] = w;
  if (!HR) sc.sq[x] = swe_sqrt(h);
}

/* one tile edge: Roe flux (interior: ApplyInteriorFlux[HR]; boundary: ApplyBoundaryFlux), flux * length stored over the edge's
 * geometry in the stage */
template <int ME, bool HR, bool COURANT, class Stage, class Scratch>
__device__ __forceinline__ void split_edge(Stage &st, Scratch &sc, const TiledArgs &p, int e, int eo, double tiny, double ha2, double dt, double &cmax,
                                           int &carg) {
  const SweArgs &a = p.a;
  const double g = RDYB_GRAVITY;
  const uint32_t lr = st.elr[e];
  const int l = lr & 0xffffu, r = lr >> 16;
  const double sn = st.esn[e], cn = st.ecn[e], len = st.elen[e];
  double hl = st.h[l];
  RoeFlux f = {0.0, 0.0, 0.0, 0.0};
  double k1l = 0.0, k2l = 0.0, k1r = 0.0, k2r = 0.0;
  if (r != 0xffff) {
    double hr = st.h[r];
    if (!(hl < tiny && hr < tiny)) { /* src/swe/swe_petsc.c:266 */
      const double ul = sc.u[l], vl = sc.v[l], ur = sc.u[r], vr = sc.v[r];
      bool do_flux = true;
      double sql, sqr;
      if (HR) { /* src/swe/swe_petsc.c:1062-1127 */
        const double zl = st.z[l], zr = st.z[r], zmax = fmax(zl, zr);
        const double hls = fmax(0.0, (hl + zl) - zmax), hrs = fmax(0.0, (hr + zr) - zmax);
        const double corl = 0.5 * g * (hl * hl - hls * hls) * len, corr = 0.5 * g * (hr * hr - hrs * hrs) * len;
        k1l = corl * cn; k2l = corl * sn; k1r = corr * cn; k2r = corr * sn;
        hl = hls; hr = hrs;
        do_flux = (hls > tiny || hrs > tiny);
        sql = swe_sqrt(hl); sqr = swe_sqrt(hr);
      } else {
        sql = sc.sq[l]; sqr = sc.sq[r];
      }
      if (do_flux) {
        f = swe_roe_flux(hl, ul, vl, sql, hr, ur, vr, sqr, sn, cn);
        if (COURANT) {
          const double cnum = f.amax * len * fmax(st.ia[l], st.ia[r]);
          if (cnum > cmax) { cmax = cnum; carg = eo + e; }
        }
      }
    }
  } else {
    /* boundary edge (rare): ApplyBoundaryFlux, src/swe/swe_petsc.c:506-630 */
    const int j = p.edge_bidx[eo + e];
    const double h = hl;
    double ul, vl;
    swe_velocities<false>(h, st.hu[l], st.hv[l], tiny, ha2, ul, vl);
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
      if (COURANT) {
        const double cnum = f.amax * len * st.ia[l];
        if (cnum > cmax) { cmax = cnum; carg = eo + e; }
      }
    }
    if (a.bflux) {
      a.bflux[j] = f.f0; a.bflux[a.nbe + j] = f.f1; a.bflux[2 * a.nbe + j] = f.f2;
      a.bacc[j] += dt * f.f0; a.bacc[a.nbe + j] += dt * f.f1; a.bacc[2 * a.nbe + j] += dt * f.f2;
    }
  }
  const double fl1 = f.f1 * len, fl2 = f.f2 * len;
  st.esn[e] = f.f0 * len; st.ecn[e] = fl1 + k1l; st.elen[e] = fl2 + k2l; /* in place of the edge geometry */
  if (HR) { st.f1r[e] = fl1 + k1r; st.f2r[e] = fl2 + k2r; }
}

/* one tile cell: flux sum in the reference's slot order (branch-free), bed slope, friction, rain, stage update */
template <int ME, bool HR, int SRC, bool EULER, class Stage, class Scratch>
__device__ __forceinline__ void split_cell(const Stage &st, const Scratch &sc, const SweArgs &a, int t, int i, int width, double tiny, double ha2, double dt) {
  const double g = RDYB_GRAVITY;
  const double h = st.h[t], hu = st.hu[t], hv = st.hv[t], ia = st.ia[t];
  const int ia_hi = __double2hiint(ia), ia_lo = __double2loint(ia);
  double F0 = 0.0, F1 = 0.0, F2 = 0.0;
#pragma unroll
  for (int k = 0; k < RDYB_ELL_WIDTH; ++k) {
    if (k < width) {
      const uint32_t ce = st.ce[k][t];
      const int e = min((int)(ce & 0x7fffu), ME - 1);  /* empty slot (0xffff) -> the zero flux entry */
      const int flip = (int)((~ce & 0x8000u) << 16);   /* left cell: -1/area, right cell: +1/area */
      const double sgn = __hiloint2double(ia_hi ^ flip, ia_lo);
      F0 = fma(st.esn[e], sgn, F0);
      if (HR) {
        const bool right = (ce & 0x8000u) != 0;
        F1 = fma(right ? st.f1r[e] : st.ecn[e], sgn, F1);
        F2 = fma(right ? st.f2r[e] : st.elen[e], sgn, F2);
      } else {
        F1 = fma(st.ecn[e], sgn, F1); F2 = fma(st.elen[e], sgn, F2);
      }
    }
  }
  if (a.fdiv) { a.fdiv[i] = F0; a.fdiv[a.stride + i] = F1; a.fdiv[2 * a.stride + i] = F2; }
  double bedx = 0.0, bedy = 0.0, tbx = 0.0, tby = 0.0;
  if (!HR) { bedx = st.dzdx[t] * g * h; bedy = st.dzdy[t] * g * h; }
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

template <int TC, int HM, int ME, int NT, bool HR, int SRC, bool EULER, bool COURANT>
__global__ void __launch_bounds__(NT, 2) swe_tiled_split_kernel(const FixedArgs q) {
  static_assert(TC % 32 == 0 && NT % 32 == 0 && NT > TC, "the cell group and the primitive-state group are whole warps");
  static_assert(HM <= RDYB_SPLIT_REC_HALO, "a tile record holds the halo list");
  extern __shared__ __align__(128) char smem_raw[];
  using Smem = SplitSmem<TC, HM, ME, HR>;
  Smem &S = *reinterpret_cast<Smem *>(smem_raw);
  const TiledArgs &p = q.t;
  const SweArgs &a = p.a;
  static_assert(NT == 384 && RDYB_ELL_WIDTH <= 4, "split_issue_tile hands one bulk copy to each of 12 warps");
  constexpr int NP = NT - TC;     /* threads of the primitive-state group */
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int G = gridDim.x;
  const int first = q.tile_begin + blockIdx.x;
  const int K = (q.tile_end - first + G - 1) / G; /* tiles of this CTA: first, first + G, ... */
  if (tid == 0) {
    for (int s = 0; s < 3; ++s) {
      mbar_init(&S.full_bar[s], 1 + RDYB_SPLIT_GATHER_THREADS);
      /* the zero flux entry that empty cell-edge slots point at (never touched by the bulk copies: nedge <= ME - 2) */
      S.st[s].esn[ME - 1] = 0.0; S.st[s].ecn[ME - 1] = 0.0; S.st[s].elen[ME - 1] = 0.0;
      if (HR) { S.st[s].f1r[ME - 1] = 0.0; S.st[s].f2r[ME - 1] = 0.0; }
    }
    for (int s = 0; s < 4; ++s) mbar_init(&S.rec_bar[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const size_t rec_bytes = RDYB_SPLIT_REC_INTS * sizeof(int32_t);
  if (tid == 0) {
    for (int j = 0; j < 3 && j < K; ++j) {
      mbar_expect_tx(&S.rec_bar[j], (uint32_t)rec_bytes);
      bulk_g2s(S.rec[j], p.tile_rec + (size_t)(first + j * G) * RDYB_SPLIT_REC_INTS, (uint32_t)rec_bytes, &S.rec_bar[j]);
    }
  }
  split_issue_tile<TC, HM, ME, HR>(S, p, 0, tid);
  if (K > 1) split_issue_tile<TC, HM, ME, HR>(S, p, 1, tid);
  const double tiny = a.tiny_h, ha2 = a.ha2, dt = a.dt;
  const int width = a.width;
  double cmax = 0.0;
  int carg = -1;
  /* primitive states of the first tile: everybody */
  mbar_wait(&S.full_bar[0], 0);
  for (int x = tid; x < TC + S.st[0].desc[5]; x += NT) split_prim<HR>(S.st[0], S.sc, x, tiny, ha2);
  __syncthreads();

  for (int k = 0; k < K; ++k) {
    auto &st = S.st[k % 3];
    /* stage (k+2) % 3 held tile k-1, which every thread left at the barrier that ended iteration k-1 */
    if (k + 2 < K) split_issue_tile<TC, HM, ME, HR>(S, p, k + 2, tid);
    if (tid == NT - 32 && k + 3 < K) { /* record of tile k+3 into the slot of record k-1 (read during iteration k-3) */
      uint64_t *rb = &S.rec_bar[(k + 3) & 3];
      mbar_expect_tx(rb, (uint32_t)rec_bytes);
      bulk_g2s(S.rec[(k + 3) & 3], p.tile_rec + (size_t)(first + (k + 3) * G) * RDYB_SPLIT_REC_INTS, (uint32_t)rec_bytes, rb);
    }
    /* ---- region A: one flux per tile edge ---- */
    mbar_wait(&S.full_bar[k % 3], (k / 3) & 1); /* completed long ago (the primitive-state group waited for it); makes the copies visible to this thread */
    const int cb = st.desc[0], ncell = st.desc[1], eo = st.desc[2], nedge = st.desc[3];
    for (int e = tid; e < nedge; e += NT) split_edge<ME, HR, COURANT>(st, S.sc, p, e, eo, tiny, ha2, dt, cmax, carg);
    __syncthreads();
    /* ---- region B: cells of this tile | primitive states of the next tile ---- */
    if (tid < TC) {
      if (tid < ncell) split_cell<ME, HR, SRC, EULER>(st, S.sc, a, tid, cb + tid, width, tiny, ha2, dt);
    } else if (k + 1 < K) {
      auto &sn = S.st[(k + 1) % 3];
      mbar_wait(&S.full_bar[(k + 1) % 3], ((k + 1) / 3) & 1);
      const int nx = TC + sn.desc[5];
      for (int x = tid - TC; x < nx; x += NP) split_prim<HR>(sn, S.sc, x, tiny, ha2);
    }
    fence_proxy_async_smem(); /* this thread's writes into the stage (fluxes, in place) before the bulk copies that will reuse it */
    __syncthreads();
  }

  if (COURANT && a.cmax) {
    cmax *= dt;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const double ov = __shfl_down_sync(0xffffffffu, cmax, off);
      const int oa = __shfl_down_sync(0xffffffffu, carg, off);
      if (ov > cmax || (ov == cmax && oa >= 0 && (carg < 0 || oa < carg))) { cmax = ov; carg = oa; }
    }
    if (lane == 0) { S.wval[warp] = cmax; S.wargs[warp] = carg; }
    __syncthreads();
    if (tid == 0) {
      for (int w = 1; w < NT / 32; ++w)
        if (S.wval[w] > cmax || (S.wval[w] == cmax && S.wargs[w] >= 0 && (carg < 0 || S.wargs[w] < carg))) { cmax = S.wval[w]; carg = S.wargs[w]; }
      const int slot = q.tile_begin + blockIdx.x; /* tile numbers of different launches do not overlap, so neither do the slots */
      const long long arg = carg;
      if (cmax > a.cmax[slot] || (cmax == a.cmax[slot] && arg >= 0 && (a.carg[slot] < 0 || arg < a.carg[slot]))) { a.cmax[slot] = cmax; a.carg[slot] = arg; }
    }
  }
}

/* ---------------- host side ---------------- */
template <int TC, int HM, int ME, int NT, bool HR, int SRC, bool EULER, bool COURANT>
static int split_launch_t(const FixedArgs &q, int ntiles, cudaStream_t stream) {
  static PersistentCache cache = {};
  auto kern = swe_tiled_split_kernel<TC, HM, ME, NT, HR, SRC, EULER, COURANT>;
  const size_t smem = sizeof(SplitSmem<TC, HM, ME, HR>);
  int grid = 0;
  RDYB_CALL(persistent_grid(kern, NT, smem, ntiles, cache, "split tile kernel", &grid));
  kern<<<grid, NT, smem, stream>>>(q);
  return RDYB_SUCCESS;
}

template <int TC, int HM, int ME, int NT, bool HR>
static int split_launch_cfg(const FixedArgs &q, int nt, bool xq, bool euler, bool courant, cudaStream_t stream) {
#define SL(SRCv)                                                                                              \
  (euler ? (courant ? split_launch_t<TC, HM, ME, NT, HR, SRCv, true, true>(q, nt, stream)                     \
                    : split_launch_t<TC, HM, ME, NT, HR, SRCv, true, false>(q, nt, stream))                   \
         : (courant ? split_launch_t<TC, HM, ME, NT, HR, SRCv, false, true>(q, nt, stream)                    \
                    : split_launch_t<TC, HM, ME, NT, HR, SRCv, false, false>(q, nt, stream)))
  return xq ? SL(1) : SL(0);
#undef SL
}

/* the compiled shape: {tile cells, halo capacity, edge capacity (last entry = the zero flux), threads}: triangle meshes, 224-cell
 * tiles (~370 edges: one per thread) */
#define RDYB_SPLIT_S 224, 80, 400, 384

/* K6 runs tiles of 224 cells with <= 80 halo cells and <= 398 edges, in index order, without momentum sources */
static inline bool tiled_split_applicable(const TiledDev &d, int tc, int mh, int me, const SweArgs &a) {
  if (a.s1 || a.s2 || !d.tile_rec || d.order) return false;
  return tc == 224 && mh <= 80 && me <= 398;
}

static inline int tiled_split_launch(const TiledDev &d, const RdybTiledLayout &T, const RdybConfig &cfg, const SweArgs &a, int32_t n_interior,
                                     bool euler, bool courant, cudaStream_t stream) {
  FixedArgs q;
  TiledArgs &p = q.t;
  p.a = a;
  p.tile_desc = d.tile_desc; p.tile_cell_begin = d.tile_cell_begin; p.halo_offsets = d.halo_offsets; p.halo_cells = d.halo_cells;
  p.edge_offsets = d.edge_offsets; p.edge_bidx = d.edge_bidx; p.edge_lr = d.edge_lr; p.edge_sn = d.edge_sn; p.edge_cn = d.edge_cn;
  p.edge_len = d.edge_len; p.cell_edges = d.cell_edges; p.tile_rec = d.tile_rec;
  p.tile0 = 0; p.max_cells = T.tile_cells; p.max_staged = 0; p.max_edges = 0;
  p.tile_order = nullptr;
  q.tile_begin = a.begin < n_interior ? 0 : T.num_interior_tiles;
  q.tile_end = (a.end <= n_interior && a.begin < n_interior) ? T.num_interior_tiles : T.num_tiles;
  if (q.tile_end <= q.tile_begin) return RDYB_SUCCESS;
  const bool hr = cfg.well_balancing == RDYB_WELL_BALANCING_HR;
  const bool xq = cfg.source_method == RDYB_SOURCE_IMPLICIT_XQ2018;
  const int nt = q.tile_end - q.tile_begin;
  return hr ? split_launch_cfg<RDYB_SPLIT_S, true>(q, nt, xq, euler, courant, stream) : split_launch_cfg<RDYB_SPLIT_S, false>(q, nt, xq, euler, courant, stream);
}
#endif
