/* swe_warp.cuh — K5: persistent TMA-staged tile kernel in which every consumer WARP owns a 32-cell mini-tile.
 *
 * K3/K4 split a tile's work into three CTA-wide phases separated by named barriers; their ncu captures
 * (profiles/) show ~17 % of all warp-cycles parked on those barriers (the edge phase has 1.6 edges per thread, i.e. two
 * passes with the second one 60 % full) and long dependent fp64 chains exposed by the low warp count.  K5 keeps the
 * producer warp, the cp.async.bulk (TMA, SASS UBLKCP) staging and the mbarrier pipeline of K4, but cuts every tile
 * into 32-cell mini-tiles (device_layout.h, RdybWarpLayout): a consumer warp
 *   (1) computes the primitive state of its 32 cells and of the <= 32 outside cells its edges touch,
 *   (2) evaluates its ~60 edges two at a time per lane (independent Roe fluxes interleave -> ILP 2; the rare
 *       critical-flow fix and boundary edges take a separate slow path chosen by a warp vote),
 *   (3) sums each cell's fluxes branch-free, applies the sources and writes u + dt F,
 * with only __syncwarp() in between.  Warps of a CTA drift apart, so fp64-heavy and shared-memory-heavy phases of
 * different warps overlap.  Edges between two mini-tiles are evaluated by both (identical inputs and orientation ->
 * identical bits -> exact conservation), 1.8 instead of 1.6 flux evaluations per triangle.
 *
 * Replaces (fused): ApplyInteriorFlux src/swe/swe_petsc.c:215-316, ApplyInteriorFluxHR :1000-1161, ApplyBoundaryFlux
 * :506-630, ApplySourceSemiImplicit :704-804, ApplySourceImplicitXQ2018 :816-932 and the TSEULER update.
 */
#ifndef RDYB_SWE_WARP_CUH
#define RDYB_SWE_WARP_CUH
#include <type_traits>

#include "swe_tiled_fixed.cuh"

#ifndef RDYB_WARP_ILP
#define RDYB_WARP_ILP 1
#endif

struct WarpDev {
  int32_t *tile_desc, *halo_cells, *edge_bidx;
  uint16_t *wdesc, *whalo, *edge_lr;
  double *edge_sn, *edge_cn, *edge_len;
  uint8_t *cell_edges;
};

struct WarpArgs {
  SweArgs a;
  const int32_t *__restrict__ tile_desc, *__restrict__ halo_cells, *__restrict__ edge_bidx;
  const uint16_t *__restrict__ wdesc, *__restrict__ whalo, *__restrict__ edge_lr;
  const double *__restrict__ edge_sn, *__restrict__ edge_cn, *__restrict__ edge_len;
  const uint8_t *__restrict__ cell_edges;
  int32_t tile_begin, tile_end;
};

/* one pipeline stage: everything the producer copies for one tile */
template <int TC, int HM, int ME, int WH, bool HR>
struct alignas(128) WarpStage {
  static constexpr int CS = TC + HM;
  double h[CS], hu[CS], hv[CS], ia[CS];
  double z[HR ? CS : 2];
  double dzdx[HR ? 2 : TC], dzdy[HR ? 2 : TC];
  double mann[TC], s0[TC];
  double esn[ME], ecn[ME], elen[ME];
  uint16_t elr[ME];
  uint16_t whalo[WH];
  uint16_t wdesc[TC / 32 * 4];
  uint8_t ce[RDYB_ELL_WIDTH][TC];
  int32_t desc[8];
};

/* per-warp scratch: primitive state of the 32 own cells + up to 32 outside cells, and up to EC edge fluxes
 * (entry EC-1 is the zero flux that empty cell-edge slots point at) */
template <int EC, bool HR>
struct alignas(16) WarpScratch {
  double h[64], u[64], v[64], ia[64];
  double sq[HR ? 2 : 64], z[HR ? 64 : 2];
  double f0[EC], f1[EC], f2[EC];
  double f1r[HR ? EC : 2], f2r[HR ? EC : 2]; /* HR: momentum flux + pressure correction as seen by the right cell (f1, f2: by the left cell) */
};

template <int TC, int HM, int ME, int WH, int EC, int NS, int NWC, bool HR>
struct WarpSmem {
  WarpStage<TC, HM, ME, WH, HR> st[NS]; /* NS-deep pipeline of staged tiles */
  WarpScratch<EC, HR> ws[NWC];          /* one scratch per consumer warp */
  uint64_t full_bar[NS], empty_bar[NS];
  double wval[NWC];
  int wargs[NWC];
  int claim;                            /* next unclaimed (tile, mini-tile) work item of this CTA */
};

template <int TC, int HM, int ME, int WH, int EC, int NS, int NWC, int MINB, bool HR, int SRC, bool EULER, bool COURANT>
__global__ void __launch_bounds__(NWC * 32 + 32, MINB) swe_warp_kernel(const WarpArgs p) {
  extern __shared__ __align__(128) char smem_raw[];
  using Smem = WarpSmem<TC, HM, ME, WH, EC, NS, NWC, HR>;
  Smem &S = *reinterpret_cast<Smem *>(smem_raw);
  const SweArgs &a = p.a;
  constexpr int NW = NWC;            /* consumer warps */
  constexpr int NMT = TC / 32;       /* mini-tiles per staged tile */
  constexpr int HK = (HM + 31) / 32; /* halo indices per producer lane */
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < NS; ++s) { mbar_init(&S.full_bar[s], 33); mbar_init(&S.empty_bar[s], TC); }
    S.claim = 0;
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp < NW && lane == 0) {
    auto &ws = S.ws[warp];
    ws.f0[EC - 1] = 0.0; ws.f1[EC - 1] = 0.0; ws.f2[EC - 1] = 0.0;
    if (HR) { ws.f1r[EC - 1] = 0.0; ws.f2r[EC - 1] = 0.0; }
  }
  __syncthreads();
  const int G = gridDim.x;

  if (warp == NW) {
    /* ================= producer warp ================= */
    int tile = p.tile_begin + blockIdx.x;
    const int4 zero4 = make_int4(0, 0, 0, 0);
    int4 d0 = zero4, d1 = zero4, n0 = zero4, n1 = zero4;
    if (tile < p.tile_end) {
      d0 = __ldg(reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)tile));
      d1 = __ldg(reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)tile + 4));
    }
    if (tile + G < p.tile_end) {
      n0 = __ldg(reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)(tile + G)));
      n1 = __ldg(reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)(tile + G) + 4));
    }
    int hidx[HK];
#pragma unroll
    for (int k = 0; k < HK; ++k) { const int t = lane + 32 * k; hidx[k] = t < d1.y ? __ldg(p.halo_cells + d1.x + t) : 0; }
    for (int it = 0; tile < p.tile_end; tile += G, ++it) {
      const int s = it % NS;
      /* prefetch: descriptor two tiles ahead, halo index list one tile ahead */
      int4 nn0 = zero4, nn1 = zero4;
      if (tile + 2 * G < p.tile_end) {
        nn0 = __ldg(reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)(tile + 2 * G)));
        nn1 = __ldg(reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)(tile + 2 * G) + 4));
      }
      int nh[HK];
#pragma unroll
      for (int k = 0; k < HK; ++k) { const int t = lane + 32 * k; nh[k] = t < n1.y ? __ldg(p.halo_cells + n1.x + t) : 0; }
      mbar_wait(&S.empty_bar[s], ((it / NS) & 1) ^ 1);
      auto &st = S.st[s];
      const int cb = d0.x, ncell = d0.y, eo = d0.z, nedge = d0.w, nhalo = d1.y, wo = d1.z, nwh = d1.w;
      uint64_t *bar = &S.full_bar[s];
      if (lane == 0) {
        st.desc[0] = cb; st.desc[1] = ncell; st.desc[2] = eo;
        const uint32_t bc = (uint32_t)((ncell + 1) & ~1) * 8u;  /* per-cell double slices */
        const uint32_t b8 = (uint32_t)((ncell + 15) & ~15);     /* per-cell uint8 slices */
        const uint32_t be = (uint32_t)((nedge + 1) & ~1) * 8u;  /* per-edge double slices */
        const uint32_t bl = (uint32_t)((nedge + 7) & ~7) * 2u;  /* per-edge uint16 slice */
        const uint32_t bw = (uint32_t)((nwh + 7) & ~7) * 2u;    /* mini-tile halo lists */
        constexpr uint32_t bd = NMT * 8u;                       /* mini-tile descriptors */
        constexpr int ncd = 3 + 1 + (HR ? 1 : 2) + 1 + 1;
        mbar_expect_tx(bar, bc * ncd + b8 * a.width + be * 3 + bl + bw + bd);
        bulk_g2s(st.h, a.h + cb, bc, bar);
        bulk_g2s(st.hu, a.hu + cb, bc, bar);
        bulk_g2s(st.hv, a.hv + cb, bc, bar);
        bulk_g2s(st.ia, a.inv_area + cb, bc, bar);
        if (HR) bulk_g2s(st.z, a.zc + cb, bc, bar);
        else { bulk_g2s(st.dzdx, a.dzdx + cb, bc, bar); bulk_g2s(st.dzdy, a.dzdy + cb, bc, bar); }
        bulk_g2s(st.mann, a.mann + cb, bc, bar);
        bulk_g2s(st.s0, a.s0 + cb, bc, bar);
        for (int k = 0; k < a.width; ++k) bulk_g2s(st.ce[k], p.cell_edges + (size_t)k * a.stride + cb, b8, bar);
        bulk_g2s(st.esn, p.edge_sn + eo, be, bar);
        bulk_g2s(st.ecn, p.edge_cn + eo, be, bar);
        bulk_g2s(st.elen, p.edge_len + eo, be, bar);
        bulk_g2s(st.elr, p.edge_lr + eo, bl, bar);
        if (bw) bulk_g2s(st.whalo, p.whalo + wo, bw, bar);
        bulk_g2s(st.wdesc, p.wdesc + (size_t)tile * (NMT * 4), bd, bar);
      }
      /* halo gather: 8-byte async copies straight into the staged arrays at local index TC + t */
#pragma unroll
      for (int k = 0; k < HK; ++k) {
        const int t = lane + 32 * k;
        if (t < nhalo) {
          const int c = hidx[k];
          cp_async8(st.h + TC + t, a.h + c);
          cp_async8(st.hu + TC + t, a.hu + c);
          cp_async8(st.hv + TC + t, a.hv + c);
          cp_async8(st.ia + TC + t, a.inv_area + c);
          if constexpr (HR) cp_async8(st.z + TC + t, a.zc + c);
        }
      }
      cp_async_arrive_noinc(bar);
      d0 = n0; d1 = n1; n0 = nn0; n1 = nn1;
#pragma unroll
      for (int k = 0; k < HK; ++k) hidx[k] = nh[k];
    }
    return;
  }

  /* ================= consumer warps: claim (tile, mini-tile) work items in order, one 32-cell mini-tile at a time ================= */
  auto &ws = S.ws[warp];
  const double g = RDYB_GRAVITY, tiny = a.tiny_h, ha2 = a.ha2, dt = a.dt;
  const int width = a.width;
  double cmax = 0.0;
  int carg = -1;
  const int my_tiles = (p.tile_end - p.tile_begin - (int)blockIdx.x + G - 1) / G; /* tiles this CTA walks */
  const int nitems = my_tiles * NMT;
  for (;;) {
    int item = 0;
    if (lane == 0) item = atomicAdd(&S.claim, 1);
    item = __shfl_sync(0xffffffffu, item, 0);
    if (item >= nitems) break;
    const int it = item / NMT, mt = item - it * NMT;
    const int s = it % NS;
    mbar_wait(&S.full_bar[s], (it / NS) & 1);
    auto &st = S.st[s];
    const int cb = st.desc[0], ncell = st.desc[1], eo = st.desc[2];
    const uint2 wdp = *reinterpret_cast<const uint2 *>(&st.wdesc[4 * mt]);
    const int ebeg = wdp.x & 0xffffu, nint = (wdp.x >> 16) & 0xffu, nbnd = wdp.x >> 24, hbeg = wdp.y & 0xffffu, nwh = wdp.y >> 16;
    const int t = 32 * mt + lane; /* this lane's cell within the tile */

    /* ---- phase 1: primitive state of the lane's own cell and of one outside cell of the mini-tile ---- */
    {
      const int xh = lane < nwh ? st.whalo[hbeg + lane] : t;
      const double h0 = st.h[t], h1 = st.h[xh];
      double u0, v0, u1, v1;
      swe_velocities<HR>(h0, st.hu[t], st.hv[t], tiny, ha2, u0, v0);
      swe_velocities<HR>(h1, st.hu[xh], st.hv[xh], tiny, ha2, u1, v1);
      ws.h[lane] = h0; ws.u[lane] = u0; ws.v[lane] = v0; ws.ia[lane] = st.ia[t];
      ws.h[32 + lane] = h1; ws.u[32 + lane] = u1; ws.v[32 + lane] = v1; ws.ia[32 + lane] = st.ia[xh];
      if (HR) { ws.z[lane] = st.z[t]; ws.z[32 + lane] = st.z[xh]; }
      else { ws.sq[lane] = swe_sqrt(h0); ws.sq[32 + lane] = swe_sqrt(h1); }
    }
    __syncwarp();

    /* ---- phase 2: the mini-tile's interior edges, two per lane and pass (branch-free main path) ---- */
    /* flux of mini-tile edge i (times its length), for the left / right cell; ok = the edge carries a Roe flux */
    auto edge_flux = [&](auto fix_tag, int i, bool valid, double &o0, double &o1l, double &o2l, double &o1r, double &o2r, double &cnum, bool &needfix) {
      constexpr bool FIX = decltype(fix_tag)::value;
      const uint32_t lr = st.elr[ebeg + i];
      const int l = lr & 0xffu, r = lr >> 8;
      const double sn = st.esn[ebeg + i], cn = st.ecn[ebeg + i], len = st.elen[ebeg + i];
      double hl = ws.h[l], hr = ws.h[r];
      const double ul = ws.u[l], vl = ws.v[l], ur = ws.u[r], vr = ws.v[r];
      bool ok = valid && !(hl < tiny && hr < tiny);
      double k1l = 0.0, k2l = 0.0, k1r = 0.0, k2r = 0.0, sql, sqr;
      if (HR) {
        const double zl = ws.z[l], zr = ws.z[r], zmax = fmax(zl, zr);
        const double hls = fmax(0.0, (hl + zl) - zmax), hrs = fmax(0.0, (hr + zr) - zmax);
        const double corl = 0.5 * g * (hl * hl - hls * hls) * len, corr = 0.5 * g * (hr * hr - hrs * hrs) * len;
        k1l = ok ? corl * cn : 0.0; k2l = ok ? corl * sn : 0.0; k1r = ok ? corr * cn : 0.0; k2r = ok ? corr * sn : 0.0;
        hl = hls; hr = hrs;
        ok = ok && (hls > tiny || hrs > tiny);
        sql = swe_sqrt(hl); sqr = swe_sqrt(hr);
      } else {
        sql = ws.sq[l]; sqr = ws.sq[r];
      }
      bool nf;
      const RoeFlux f = swe_roe_flux_t<FIX>(hl, ul, vl, sql, hr, ur, vr, sqr, sn, cn, nf);
      needfix = nf && ok;
      const double f0 = ok ? f.f0 * len : 0.0, f1 = ok ? f.f1 * len : 0.0, f2 = ok ? f.f2 * len : 0.0;
      o0 = f0; o1l = f1 + k1l; o2l = f2 + k2l; o1r = f1 + k1r; o2r = f2 + k2r;
      cnum = ok ? f.amax * len * fmax(ws.ia[l], ws.ia[r]) : 0.0;
    };
#if RDYB_WARP_ILP == 2
    for (int base = 0; base < nint; base += 64) {
      const int ea = base + lane, eb = ea + 32;
      const bool va = ea < nint, vb = eb < nint;
      const int ia_ = va ? ea : 0, ib_ = vb ? eb : 0;
      double a0, a1l, a2l, a1r, a2r, ac, b0, b1l, b2l, b1r, b2r, bc;
      bool nfa, nfb;
      edge_flux(std::false_type{}, ia_, va, a0, a1l, a2l, a1r, a2r, ac, nfa);
      edge_flux(std::false_type{}, ib_, vb, b0, b1l, b2l, b1r, b2r, bc, nfb);
      if (__any_sync(0xffffffffu, nfa || nfb)) { /* critical-flow fix needed somewhere in the warp (rare): redo those edges */
        bool dummy;
        if (nfa) edge_flux(std::true_type{}, ia_, va, a0, a1l, a2l, a1r, a2r, ac, dummy);
        if (nfb) edge_flux(std::true_type{}, ib_, vb, b0, b1l, b2l, b1r, b2r, bc, dummy);
      }
      if (va) {
        ws.f0[ea] = a0; ws.f1[ea] = a1l; ws.f2[ea] = a2l;
        if (HR) { ws.f1r[ea] = a1r; ws.f2r[ea] = a2r; }
      }
      if (vb) {
        ws.f0[eb] = b0; ws.f1[eb] = b1l; ws.f2[eb] = b2l;
        if (HR) { ws.f1r[eb] = b1r; ws.f2r[eb] = b2r; }
      }
      if (COURANT) {
        if (ac > cmax) { cmax = ac; carg = eo + ebeg + ea; }
        if (bc > cmax) { cmax = bc; carg = eo + ebeg + eb; }
      }
    }
#else
    for (int ea = lane; ea < nint; ea += 32) {
      double a0, a1l, a2l, a1r, a2r, ac;
      bool nfa;
      edge_flux(std::false_type{}, ea, true, a0, a1l, a2l, a1r, a2r, ac, nfa);
      if (__any_sync(__activemask(), nfa)) { /* critical-flow fix needed somewhere in the warp (rare): redo those edges */
        bool dummy;
        if (nfa) edge_flux(std::true_type{}, ea, true, a0, a1l, a2l, a1r, a2r, ac, dummy);
      }
      ws.f0[ea] = a0; ws.f1[ea] = a1l; ws.f2[ea] = a2l;
      if (HR) { ws.f1r[ea] = a1r; ws.f2r[ea] = a2r; }
      if (COURANT) {
        if (ac > cmax) { cmax = ac; carg = eo + ebeg + ea; }
      }
    }
#endif
    /* boundary edges of the mini-tile (rare): ApplyBoundaryFlux, src/swe/swe_petsc.c:506-630 */
    for (int e = nint + lane; e < nint + nbnd; e += 32) {
      const int l = st.elr[ebeg + e] & 0xffu, xl = 32 * mt + l;
      const double sn = st.esn[ebeg + e], cn = st.ecn[ebeg + e], len = st.elen[ebeg + e];
      const int j = p.edge_bidx[eo + ebeg + e];
      const double h = ws.h[l];
      double ul, vl;
      swe_velocities<false>(h, st.hu[xl], st.hv[xl], tiny, ha2, ul, vl);
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
      RoeFlux f = {0.0, 0.0, 0.0, 0.0};
      if (ok && !(h < tiny && hr < tiny)) {
        f = swe_roe_flux(h, ul, vl, swe_sqrt(h), hr, ur, vr, swe_sqrt(hr), sn, cn);
        if (COURANT) {
          const double cnum = f.amax * len * ws.ia[l];
          if (cnum > cmax) { cmax = cnum; carg = eo + ebeg + e; }
        }
      }
      if (a.bflux) {
        a.bflux[j] = f.f0; a.bflux[a.nbe + j] = f.f1; a.bflux[2 * a.nbe + j] = f.f2;
        a.bacc[j] += dt * f.f0; a.bacc[a.nbe + j] += dt * f.f1; a.bacc[2 * a.nbe + j] += dt * f.f2;
      }
      ws.f0[e] = f.f0 * len; ws.f1[e] = f.f1 * len; ws.f2[e] = f.f2 * len;
      if (HR) { ws.f1r[e] = f.f1 * len; ws.f2r[e] = f.f2 * len; }
    }
    __syncwarp();

    /* ---- phase 3: flux sum of the lane's cell (branch-free), sources, update ---- */
    if (t < ncell) {
      const int i = cb + t;
      const double h = st.h[t], hu = st.hu[t], hv = st.hv[t], ia = st.ia[t];
      const int ia_hi = __double2hiint(ia), ia_lo = __double2loint(ia);
      double F0 = 0.0, F1 = 0.0, F2 = 0.0;
#pragma unroll
      for (int k = 0; k < RDYB_ELL_WIDTH; ++k) {
        if (k < width) {
          const uint32_t ce = st.ce[k][t];
          const int e = min((int)(ce & 0x7fu), EC - 1);   /* empty slot (0xff) -> the zero flux entry */
          const int flip = (int)((~ce & 0x80u) << 24);    /* left cell: -1/area, right cell: +1/area */
          const double sgn = __hiloint2double(ia_hi ^ flip, ia_lo);
          F0 = fma(ws.f0[e], sgn, F0);
          if (HR) {
            const bool right = (ce & 0x80u) != 0;
            F1 = fma(right ? ws.f1r[e] : ws.f1[e], sgn, F1);
            F2 = fma(right ? ws.f2r[e] : ws.f2[e], sgn, F2);
          } else {
            F1 = fma(ws.f1[e], sgn, F1); F2 = fma(ws.f2[e], sgn, F2);
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
    __syncwarp(); /* the warp scratch is rewritten by the next tile's phase 1 */
    mbar_arrive(&S.empty_bar[s]); /* this thread is done reading stage s */
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
    asm volatile("bar.sync 1, %0;" ::"n"(NWC * 32) : "memory");
    if (threadIdx.x == 0) {
      for (int w = 1; w < NW; ++w)
        if (S.wval[w] > cmax || (S.wval[w] == cmax && S.wargs[w] >= 0 && (carg < 0 || S.wargs[w] < carg))) { cmax = S.wval[w]; carg = S.wargs[w]; }
      const int slot = p.tile_begin + blockIdx.x; /* tile numbers of different launches do not overlap, so neither do the slots */
      const long long arg = carg;
      if (cmax > a.cmax[slot] || (cmax == a.cmax[slot] && arg >= 0 && (a.carg[slot] < 0 || arg < a.carg[slot]))) { a.cmax[slot] = cmax; a.carg[slot] = arg; }
    }
  }
}

/* ---------------- host side ---------------- */
static inline void warp_dev_free(WarpDev *d) {
  void *ptrs[] = {d->tile_desc, d->halo_cells, d->edge_bidx, d->wdesc, d->whalo, d->edge_lr, d->edge_sn, d->edge_cn, d->edge_len, d->cell_edges};
  for (void *q : ptrs)
    if (q) cudaFree(q);
  memset(d, 0, sizeof(*d));
}

template <typename T>
static int warp_upload(T **dst, const T *src, size_t n) {
  cudaError_t e = cudaMalloc((void **)dst, sizeof(T) * (n > 0 ? n : 1));
  if (e != cudaSuccess) return rdyb_fail(RDYB_ERR_CUDA, "warp layout upload: cudaMalloc of %zu bytes failed (%s)", sizeof(T) * n, cudaGetErrorString(e));
  if (n) {
    e = cudaMemcpy(*dst, src, sizeof(T) * n, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) return rdyb_fail(RDYB_ERR_CUDA, "warp layout upload: copy failed (%s)", cudaGetErrorString(e));
  }
  return RDYB_SUCCESS;
}

static inline int warp_dev_upload(const RdybWarpLayout *W, const RdybLayout *L, WarpDev *d) {
  const size_t nt = (size_t)W->num_tiles, ne = (size_t)W->total_edges;
  RDYB_CALL(warp_upload(&d->tile_desc, W->tile_desc, 8 * nt));
  RDYB_CALL(warp_upload(&d->halo_cells, W->halo_cells, (size_t)W->total_halo));
  RDYB_CALL(warp_upload(&d->wdesc, W->wdesc, 4 * (size_t)(W->tile_cells / RDYB_WARP_CELLS) * nt + 64));
  RDYB_CALL(warp_upload(&d->whalo, W->whalo, (size_t)W->total_whalo));
  RDYB_CALL(warp_upload(&d->edge_lr, W->edge_lr, ne));
  RDYB_CALL(warp_upload(&d->edge_bidx, W->edge_bidx, ne));
  RDYB_CALL(warp_upload(&d->edge_sn, W->edge_sn, ne));
  RDYB_CALL(warp_upload(&d->edge_cn, W->edge_cn, ne));
  RDYB_CALL(warp_upload(&d->edge_len, W->edge_len, ne));
  RDYB_CALL(warp_upload(&d->cell_edges, W->cell_edges, (size_t)RDYB_ELL_WIDTH * (size_t)L->stride));
  return RDYB_SUCCESS;
}

template <int TC, int HM, int ME, int WH, int EC, int NS, int NWC, int MINB, bool HR, int SRC, bool EULER, bool COURANT>
static int warp_launch_t(const WarpArgs &q, int ntiles, cudaStream_t stream) {
  static PersistentCache cache = {};
  using Smem = WarpSmem<TC, HM, ME, WH, EC, NS, NWC, HR>;
  static_assert(ME % 8 == 0 && WH % 8 == 0 && TC % 64 == 0 && HM % 2 == 0, "bulk-copy alignment of the staged slices");
  auto kern = swe_warp_kernel<TC, HM, ME, WH, EC, NS, NWC, MINB, HR, SRC, EULER, COURANT>;
  const size_t smem = sizeof(Smem);
  int grid = 0;
  RDYB_CALL(persistent_grid(kern, NWC * 32 + 32, smem, ntiles, cache, "warp tile kernel", &grid));
  kern<<<grid, NWC * 32 + 32, smem, stream>>>(q);
  return RDYB_SUCCESS;
}

template <int TC, int HM, int ME, int WH, int EC, int NS, int NWC, int MINB, bool HR>
static int warp_launch_cfg(const WarpArgs &q, int nt, bool xq, bool euler, bool courant, cudaStream_t stream) {
#define WLC(SRCv)                                                                                                        \
  (euler ? (courant ? warp_launch_t<TC, HM, ME, WH, EC, NS, NWC, MINB, HR, SRCv, true, true>(q, nt, stream)                        \
                    : warp_launch_t<TC, HM, ME, WH, EC, NS, NWC, MINB, HR, SRCv, true, false>(q, nt, stream))                      \
         : (courant ? warp_launch_t<TC, HM, ME, WH, EC, NS, NWC, MINB, HR, SRCv, false, true>(q, nt, stream)                       \
                    : warp_launch_t<TC, HM, ME, WH, EC, NS, NWC, MINB, HR, SRCv, false, false>(q, nt, stream)))
  return xq ? WLC(1) : WLC(0);
#undef WLC
}

/* compiled shapes: {tile cells, tile halo capacity, tile edge capacity, tile mini-halo capacity, mini-tile edge capacity,
 * pipeline stages, consumer warps, CTAs/SM} */
#define RDYB_WARP_SHAPE_256 256, 96, 544, 256, 96, 2, 8, 2
#define RDYB_WARP_SHAPE_128 128, 64, 288, 128, 80, 4, 8, 2

static inline bool warp_shape_fits(const RdybWarpLayout &W, int tc, int hm, int me, int wh, int ec) {
  return W.tile_cells == tc && W.max_halo <= hm && W.max_tile_edges <= me && W.max_tile_whalo <= wh && W.max_warp_edges <= ec - 1 && W.max_warp_halo <= 32;
}
static inline bool warp_applicable(const RdybWarpLayout &W, const SweArgs &a) {
  if (a.s1 || a.s2) return false; /* momentum sources: K3 */
  return warp_shape_fits(W, 256, 96, 544, 256, 96) || warp_shape_fits(W, 128, 64, 288, 128, 80);
}

static inline int warp_launch(const WarpDev &d, const RdybWarpLayout &W, const RdybConfig &cfg, const SweArgs &a, int32_t n_interior, bool euler,
                              bool courant, cudaStream_t stream) {
  WarpArgs q;
  q.a = a;
  q.tile_desc = d.tile_desc; q.halo_cells = d.halo_cells; q.edge_bidx = d.edge_bidx; q.wdesc = d.wdesc; q.whalo = d.whalo; q.edge_lr = d.edge_lr;
  q.edge_sn = d.edge_sn; q.edge_cn = d.edge_cn; q.edge_len = d.edge_len; q.cell_edges = d.cell_edges;
  q.tile_begin = a.begin < n_interior ? 0 : W.num_interior_tiles;
  q.tile_end = (a.end <= n_interior && a.begin < n_interior) ? W.num_interior_tiles : W.num_tiles;
  if (q.tile_end <= q.tile_begin) return RDYB_SUCCESS;
  const bool hr = cfg.well_balancing == RDYB_WELL_BALANCING_HR;
  const bool xq = cfg.source_method == RDYB_SOURCE_IMPLICIT_XQ2018;
  const int nt = q.tile_end - q.tile_begin;
  if (W.tile_cells == 128) {
    if (hr) return warp_launch_cfg<RDYB_WARP_SHAPE_128, true>(q, nt, xq, euler, courant, stream);
    return warp_launch_cfg<RDYB_WARP_SHAPE_128, false>(q, nt, xq, euler, courant, stream);
  }
  if (hr) return warp_launch_cfg<RDYB_WARP_SHAPE_256, true>(q, nt, xq, euler, courant, stream);
  return warp_launch_cfg<RDYB_WARP_SHAPE_256, false>(q, nt, xq, euler, courant, stream);
}
#endif
