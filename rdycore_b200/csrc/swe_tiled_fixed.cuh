/* swe_tiled_fixed.cuh — K4: the persistent warp-specialised TMA tile kernel with a COMPILE-TIME shared-memory layout.
 *
 * Same algorithm and device layout as K3 (swe_tiled_async.cuh): per tile, (1) primitive state of the staged cells,
 * (2) one Roe flux per tile edge, (3) per-cell flux sum + sources + stage update, with a producer warp that stages
 * tile n+1 through cp.async.bulk (TMA, SASS UBLKCP) + mbarriers while the consumer warps compute tile n.
 * What changes is everything that the ncu capture of K3 (profiles/r1_ncu_full_swe_tiled_async.txt) showed to be
 * overhead — 63 % of K3's issued instructions were not fp64 arithmetic:
 *   - tile capacities (TC cells, HM halo cells, ME edges) are template constants, so every staged array sits at a
 *     constant offset of one shared-memory struct and loads/stores use immediate offsets (no IMAD address chains,
 *     no kernel-parameter reloads inside the tile loop);
 *   - the per-cell flux sum is branch-free: empty slots point at a zero flux entry, the left/right sign is an XOR on
 *     the high word of 1/area;
 *   - the producer prefetches the next tiles' descriptors and halo index lists, so its dependent global loads are
 *     off the critical path of the copy pipeline;
 *   - the Courant diagnostic keeps max(amax*L/A) and multiplies by dt once at the end.
 * Meshes whose tiles do not fit the compiled capacities, momentum sources and exotic options fall back to K3.
 *
 * Replaces (fused): ApplyInteriorFlux src/swe/swe_petsc.c:215-316, ApplyInteriorFluxHR :1000-1161, ApplyBoundaryFlux
 * :506-630, ApplySourceSemiImplicit :704-804, ApplySourceImplicitXQ2018 :816-932 and the TSEULER update.
 */
#ifndef RDYB_SWE_TILED_FIXED_CUH
#define RDYB_SWE_TILED_FIXED_CUH
#include "swe_tiled_async.cuh"

/* one pipeline stage: everything the producer copies for one tile */
template <int TC, int HM, int ME, bool HR>
struct alignas(128) FixedStage {
  static constexpr int CS = TC + HM;
  double h[CS], hu[CS], hv[CS], ia[CS];
  double z[HR ? CS : 2];            /* HR: bed elevation of the staged cells */
  double dzdx[HR ? 2 : TC], dzdy[HR ? 2 : TC];
  double mann[TC], s0[TC];
  double esn[ME], ecn[ME], elen[ME];
  uint32_t elr[ME];
  uint16_t ce[RDYB_ELL_WIDTH][TC];
  int32_t desc[8];
  /* HR: the edge fluxes as seen by the RIGHT cell (momentum flux + its pressure correction; the left cell's view overwrites the
   * edge geometry in place).  Per STAGE, not per CTA: with the fused-primitive variant no CTA barrier separates the cell phase of
   * tile n from the edge phase of tile n+1, so a fast warp writes these while a slow one still sums the previous tile's. */
  double f1r[HR ? ME : 2], f2r[HR ? ME : 2];
};

/* per-CTA scratch written by the consumers */
template <int TC, int HM, int ME, bool HR, bool NOPRIM>
struct alignas(16) FixedScratch {
  static constexpr int CS = TC + HM;
  /* NOPRIM (PM != 0): the primitive state is computed per edge side in phase 2: no per-CTA scratch */
  double u[NOPRIM ? 2 : CS], v[NOPRIM ? 2 : CS], sq[(HR || NOPRIM) ? 2 : CS];
  /* the edge fluxes overwrite the edge geometry of the stage in place (esn <- f0, ecn <- f1, elen <- f2: an edge's thread reads
   * its geometry before it writes its flux).  HR: f1, f2 hold momentum flux + pressure correction as seen by the LEFT cell; the
   * RIGHT cell's view lives in the stage too (FixedStage::f1r, f2r) */
};

/* PM, where the primitive state (u, v, sqrt h) of the staged cells is computed:
 *   0  a pass of the consumer warps over the staged cells + a CTA barrier (round 1's three-pass kernel, kept for A/B runs);
 *   1  per edge side inside the edge pass (no pass, no barrier; 2 evaluations per edge instead of 1.3 per cell) - the default.
 * PM != 0 also splits the barrier between the edge pass and the cell pass (flux_bar, see phase 3).
 * (A third variant - the otherwise parked PRODUCER warp computing the primitives into the stage one tile ahead - was built and
 * measured in round 2: 0.369 ms against 0.322 ms for PM 1, the single warp cannot keep up with seven consumers; removed.) */
template <int TC, int HM, int ME, int NS, bool HR, int PM>
struct FixedSmem {
  FixedStage<TC, HM, ME, HR> st[NS]; /* NS-deep pipeline: dry tiles are load-latency bound, a deeper queue keeps loads in flight */
  FixedScratch<TC, HM, ME, HR, PM != 0> sc;
  uint64_t full_bar[NS], empty_bar[NS];
  uint64_t flux_bar; /* FUSEP: every consumer thread arrives when its edges of the tile are done (split barrier, see phase 3) */
  double wval[16];
  int wargs[16];
};

struct FixedArgs {
  TiledArgs t;
  int32_t tile_begin, tile_end;
  /* multi-GPU, peer-memory halo: tiles >= wait_tile (the frontier tiles, numbered after the interior ones) stage ghost cells, so
   * the producer warp of a CTA waits until every neighbour's counter has reached wait_epoch before it issues the first of them;
   * interior and frontier then run in one launch.  wait_tile = INT32_MAX: no wait. */
  int32_t wait_tile = 0x7fffffff, wait_tile_end = 0x7fffffff, wait_peers = 0; /* launch positions [wait_tile, wait_tile_end) */
  unsigned int wait_epoch = 0;
  const unsigned int *wait_flags = nullptr;
  int *wait_status = nullptr;
  long long wait_timeout = 40000000000ll;      /* clock64() ticks before a missing neighbour is reported in *wait_status */
  unsigned long long *wait_cycles = nullptr;   /* [0] += ticks this CTA's producer spent in the wait, [1] = max over CTAs, [2] += 1 per waiting CTA */
  /* fused halo push (Euler step only): when a frontier tile is done, its border cells are stored straight into the neighbours'
   * ghost slots of the buffer this step writes; the CTA that finishes the last frontier tile of the launch raises this rank's
   * counter in every neighbour.  Frontier tiles are then scheduled FIRST (tile_order), so the transfer overlaps the interior
   * tiles and a step is one launch.  Send list of tile t: entries [push_range[t].x, + push_range[t].y). */
  const int2 *push_range = nullptr;
  const int32_t *push_src = nullptr;    /* [entries] device cell whose new state is sent */
  double *const *push_dst = nullptr;    /* [entries] the neighbour's ghost slot (component 0) in the buffer this step writes */
  const int32_t *push_rs = nullptr;     /* [entries] the neighbour's leading dimension */
  unsigned int *push_counter = nullptr; /* tiles with sends finished so far, over all launches */
  unsigned int push_target = 0;         /* value of *push_counter when this launch's last one is done */
  unsigned int push_epoch = 0;
  unsigned int *push_flag[RDYB_MAX_PEERS] = {};
  int32_t push_peers = 0;
};

template <int NCONS>
__device__ __forceinline__ void fixed_consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(NCONS) : "memory"); }


/* the producer warp of a CTA waits here, once per launch, until every neighbour's counter has reached `epoch` (its border cells
 * of this step have landed in this rank's ghost slots).  Out of line: it runs once per launch and must not cost the tile loop
 * registers.  A neighbour that stays away longer than `timeout` ticks is reported in *status (a hung peer must not hang this GPU;
 * every host getter checks the word). */
__device__ __noinline__ void fixed_neighbour_wait(const unsigned int *flags, int peers, unsigned int epoch, long long timeout, int *status,
                                                  unsigned long long *wait_cycles, int lane) {
  long long spun = 0;
  if (lane < peers) {
    const volatile unsigned int *f = flags + lane;
    const long long t0 = clock64();
    while ((int)(*f - epoch) < 0) {
      __nanosleep(100);
      if (clock64() - t0 > timeout) { atomicExch(status, 1 + lane); break; }
    }
    spun = clock64() - t0;
  }
  if (wait_cycles) { /* how long the step waited for its neighbours: the scaling diagnostic of bench.py */
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) spun = max(spun, __shfl_xor_sync(0xffffffffu, spun, off));
    if (lane == 0) { atomicAdd(wait_cycles, (unsigned long long)spun); atomicMax(wait_cycles + 1, (unsigned long long)spun); atomicAdd(wait_cycles + 2, 1ull); }
  }
  __syncwarp();
  __threadfence_system();
}

template <int TC, int HM, int ME, int NCONS, int NS, int MINB, bool HR, int SRC, bool EULER, bool COURANT, int PM>
__global__ void __launch_bounds__(NCONS + 32, MINB) swe_tiled_fixed_kernel(const FixedArgs q) {
  extern __shared__ __align__(128) char smem_raw[];
  using Smem = FixedSmem<TC, HM, ME, NS, HR, PM>;
  constexpr bool FUSEP = PM == 1, SPLITB = PM != 0;
  Smem &S = *reinterpret_cast<Smem *>(smem_raw);
  const TiledArgs &p = q.t;
  const SweArgs &a = p.a;
  if (a.ctl && (!a.ctl->active || a.ctl->cur != a.ctl_buf)) return; /* device-side loop (rdyb_operator_advance_device): not this launch's turn */
  constexpr int HK = (HM + 31) / 32; /* halo indices per producer lane */
  if (threadIdx.x == 0) {
    for (int s = 0; s < NS; ++s) {
      mbar_init(&S.full_bar[s], 33); mbar_init(&S.empty_bar[s], NCONS);
      /* the zero flux entry that empty cell-edge slots point at (never touched by the bulk copies: nedge <= ME - 2) */
      S.st[s].esn[ME - 1] = 0.0; S.st[s].ecn[ME - 1] = 0.0; S.st[s].elen[ME - 1] = 0.0;
      if (HR) { S.st[s].f1r[ME - 1] = 0.0; S.st[s].f2r[ME - 1] = 0.0; }
    }
    mbar_init(&S.flux_bar, NCONS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int G = gridDim.x;

  if (warp == NCONS / 32) {
    /* ================= producer warp ================= */
    int tile = q.tile_begin + blockIdx.x;
    int4 d0 = make_int4(0, 0, 0, 0), n0 = d0;
    int2 d1 = make_int2(0, 0), n1 = d1;
    const int32_t *ord = p.tile_order; /* launch position -> tile, or NULL */
    if (tile < q.tile_end) {
      const int tn = ord ? __ldg(ord + tile) : tile;
      d0 = __ldg(reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)tn));
      d1 = __ldg(reinterpret_cast<const int2 *>(p.tile_desc + 8 * (size_t)tn + 4));
    }
    if (tile + G < q.tile_end) {
      const int tn = ord ? __ldg(ord + tile + G) : tile + G;
      n0 = __ldg(reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)tn));
      n1 = __ldg(reinterpret_cast<const int2 *>(p.tile_desc + 8 * (size_t)tn + 4));
    }
    int hidx[HK];
#pragma unroll
    for (int k = 0; k < HK; ++k) { const int t = lane + 32 * k; hidx[k] = t < d1.y ? __ldg(p.halo_cells + d1.x + t) : 0; }
    bool waited = false;
    for (int it = 0; tile < q.tile_end; tile += G, ++it) {
      const int s = it % NS;
      if (tile >= q.wait_tile && tile < q.wait_tile_end && !waited) { /* halo_wait_kernel, inside the step (swe_kernels.cuh) */
        fixed_neighbour_wait(q.wait_flags, q.wait_peers, q.wait_epoch, q.wait_timeout, q.wait_status, q.wait_cycles, lane);
        waited = true;
      }
      /* prefetch: descriptor two tiles ahead, halo index list one tile ahead */
      int4 nn0 = make_int4(0, 0, 0, 0);
      int2 nn1 = make_int2(0, 0);
      if (tile + 2 * G < q.tile_end) {
        const int tn = ord ? __ldg(ord + tile + 2 * G) : tile + 2 * G;
        nn0 = __ldg(reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)tn));
        nn1 = __ldg(reinterpret_cast<const int2 *>(p.tile_desc + 8 * (size_t)tn + 4));
      }
      int nh[HK];
#pragma unroll
      for (int k = 0; k < HK; ++k) { const int t = lane + 32 * k; nh[k] = t < n1.y ? __ldg(p.halo_cells + n1.x + t) : 0; }
      mbar_wait(&S.empty_bar[s], ((it / NS) & 1) ^ 1);
      auto &st = S.st[s];
      const int cb = d0.x, ncell = d0.y, eo = d0.z, nedge = d0.w, nhalo = d1.y;
      uint64_t *bar = &S.full_bar[s];
      if (lane == 0) {
        st.desc[0] = cb; st.desc[1] = ncell; st.desc[2] = eo; st.desc[3] = nedge; st.desc[5] = nhalo;
        const uint32_t bc = (uint32_t)((ncell + 1) & ~1) * 8u; /* per-cell double slices */
        const uint32_t bh = (uint32_t)((ncell + 7) & ~7) * 2u; /* per-cell uint16 slices */
        const uint32_t be = (uint32_t)((nedge + 1) & ~1) * 8u; /* per-edge double slices */
        const uint32_t bl = (uint32_t)((nedge + 3) & ~3) * 4u; /* per-edge uint32 slice */
        constexpr int ncd = 3 + 1 + (HR ? 1 : 2) + 1 + 1;
        mbar_expect_tx(bar, bc * ncd + bh * a.width + be * 3 + bl);
        bulk_g2s(st.h, a.h + cb, bc, bar);
        bulk_g2s(st.hu, a.hu + cb, bc, bar);
        bulk_g2s(st.hv, a.hv + cb, bc, bar);
        bulk_g2s(st.ia, a.inv_area + cb, bc, bar);
        if (HR) bulk_g2s(st.z, a.zc + cb, bc, bar);
        else { bulk_g2s(st.dzdx, a.dzdx + cb, bc, bar); bulk_g2s(st.dzdy, a.dzdy + cb, bc, bar); }
        bulk_g2s(st.mann, a.mann + cb, bc, bar);
        bulk_g2s(st.s0, a.s0 + cb, bc, bar);
        for (int k = 0; k < a.width; ++k) bulk_g2s(st.ce[k], p.cell_edges + (size_t)k * a.stride + cb, bh, bar);
        bulk_g2s(st.esn, p.edge_sn + eo, be, bar);
        bulk_g2s(st.ecn, p.edge_cn + eo, be, bar);
        bulk_g2s(st.elen, p.edge_len + eo, be, bar);
        bulk_g2s(st.elr, p.edge_lr + eo, bl, bar);
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

  /* ================= consumer warps ================= */
  auto &sc = S.sc;
  const double g = RDYB_GRAVITY, tiny = a.tiny_h, ha2 = a.ha2, dt = a.ctl ? a.ctl->dt_step : a.dt;
  const int tid = threadIdx.x;
  const int width = a.width;
  double cmax = 0.0;
  int carg = -1;
  int it = 0;
  for (int tile = q.tile_begin + blockIdx.x; tile < q.tile_end; tile += G, ++it) {
    const int s = it % NS;
    mbar_wait(&S.full_bar[s], (it / NS) & 1);
    auto &st = S.st[s];
    const int cb = st.desc[0], ncell = st.desc[1], eo = st.desc[2], nedge = st.desc[3], nhalo = st.desc[5];

    if (p.tile0 == 1) { /* experiment (RDYB_DEBUG_COPYONLY=1): the staging pipeline alone, no arithmetic */
      for (int t = tid; t < ncell; t += NCONS) { a.o0[cb + t] = st.h[t]; a.o1[cb + t] = st.hu[t]; a.o2[cb + t] = st.hv[t]; }
      mbar_arrive(&S.empty_bar[s]);
      continue;
    }
    /* ---- phase 1: primitive state of the staged cells: tile cells [0,TC) and halo cells [TC,TC+nhalo) ----
     * (FUSEP: folded into phase 2 - each edge derives u, v, sqrt(h) of its two sides from one MUFU seed per side, which costs
     * fewer instructions than this pass + its loads, and removes a pass, a CTA barrier and 7 KB of scratch per CTA) */
    if (PM == 0) {
      for (int x = tid; x < TC + nhalo; x += NCONS) {
        const double h = st.h[x];
        double u, w;
        swe_velocities<HR>(h, st.hu[x], st.hv[x], tiny, ha2, u, w);
        sc.u[x] = u; sc.v[x] = w;
        if (!HR) sc.sq[x] = swe_sqrt(h);
      }
      fixed_consumer_sync<NCONS>();
    }

    /* ---- phase 2: one flux per tile edge ---- */
    for (int e = tid; e < nedge; e += NCONS) {
      const uint32_t lr = st.elr[e];
      const int l = lr & 0xffffu, r = lr >> 16;
      const double sn = st.esn[e], cn = st.ecn[e], len = st.elen[e];
      double hl = st.h[l];
      RoeFlux f = {0.0, 0.0, 0.0, 0.0};
      double dl = 0.0, dr = 0.0; /* HR: h^2 - h*^2 of the two sides (the pressure corrections are formed after the Roe flux: 2 live values instead of 4) */
      if (r != 0xffff) {
        double hr = st.h[r];
        if (!(hl < tiny && hr < tiny)) {
          double ul, vl, ur, vr, sql, sqr;
          if (FUSEP) {
            if (HR) { swe_velocities<true>(hl, st.hu[l], st.hv[l], tiny, ha2, ul, vl); swe_velocities<true>(hr, st.hu[r], st.hv[r], tiny, ha2, ur, vr); }
            else { swe_edge_side(hl, st.hu[l], st.hv[l], tiny, ha2, ul, vl, sql); swe_edge_side(hr, st.hu[r], st.hv[r], tiny, ha2, ur, vr, sqr); }
          } else {
            ul = sc.u[l]; vl = sc.v[l]; ur = sc.u[r]; vr = sc.v[r];
          }
          bool do_flux = true;
          if (HR) {
            const double zl = st.z[l], zr = st.z[r], zmax = fmax(zl, zr);
            const double hls = fmax(0.0, (hl + zl) - zmax), hrs = fmax(0.0, (hr + zr) - zmax);
            dl = hl * hl - hls * hls; dr = hr * hr - hrs * hrs;
            hl = hls; hr = hrs;
            do_flux = (hls > tiny || hrs > tiny);
            sql = swe_sqrt_fast(hl); sqr = swe_sqrt_fast(hr);
          } else if (PM == 0) {
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
      if (HR) {
        const double corl = 0.5 * g * dl * len, corr = 0.5 * g * dr * len;
        st.esn[e] = f.f0 * len; st.ecn[e] = fl1 + corl * cn; st.elen[e] = fl2 + corl * sn; /* in place of the edge geometry */
        st.f1r[e] = fl1 + corr * cn; st.f2r[e] = fl2 + corr * sn;
      } else {
        st.esn[e] = f.f0 * len; st.ecn[e] = fl1; st.elen[e] = fl2; /* in place of the edge geometry */
      }
    }
    /* ---- phase 3: per-cell flux sum (branch-free), sources, update ----
     * FUSEP: the barrier between the edge pass and the cell pass is SPLIT (an mbarrier: arrive now, wait later).  Between the two a
     * thread does the part of its cell that needs no flux: loads, bed slope and the nonlinear part of the friction term
     * (swe_friction_pre: reciprocal, cube root, square root - two thirds of the cell's arithmetic).  Warps with one edge sweep
     * instead of two used to idle at the barrier; now they overlap it with this work, and only the flux sum and the update remain
     * behind the wait. */
    static_assert(TC <= NCONS, "one cell per consumer thread");
    if (SPLITB) {
      mbar_arrive(&S.flux_bar);
      const int t = tid;
      const bool has = t < ncell;
      const int i = cb + t;
      double h = 0.0, hu = 0.0, hv = 0.0, ia = 1.0, bedx = 0.0, bedy = 0.0, p0 = 0.0, p1 = 0.0, mann = 0.0, s0v = 0.0;
      if (has) {
        h = st.h[t]; hu = st.hu[t]; hv = st.hv[t]; ia = st.ia[t]; mann = st.mann[t]; s0v = st.s0[t];
        if (!HR) { bedx = st.dzdx[t] * g * h; bedy = st.dzdy[t] * g * h; }
        if (h >= tiny) swe_friction_pre<SRC>(h, hu, hv, mann, dt, p0, p1);
      }
      mbar_wait(&S.flux_bar, it & 1); /* every edge flux of the tile is in shared memory */
      if (has) {
        const int ia_hi = __double2hiint(ia), ia_lo = __double2loint(ia);
        double F0 = 0.0, F1 = 0.0, F2 = 0.0;
#pragma unroll
        for (int k = 0; k < RDYB_ELL_WIDTH; ++k) {
          if (k < width) {
            const uint32_t ce = st.ce[k][t];
            const int e = min((int)(ce & 0x7fffu), ME - 1);                  /* empty slot (0xffff) -> the zero flux entry */
            const int flip = (int)((~ce & 0x8000u) << 16);                   /* left cell: -1/area, right cell: +1/area */
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
        double tbx = 0.0, tby = 0.0;
        if (h >= tiny) swe_friction_post<SRC>(p0, p1, hu, hv, F1, F2, bedx, bedy, mann, dt, a.xq_thr, tbx, tby);
        F0 += s0v;
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
    } else {
    fixed_consumer_sync<NCONS>();
    for (int t = tid; t < ncell; t += NCONS) {
      const int i = cb + t;
      const double h = st.h[t], hu = st.hu[t], hv = st.hv[t], ia = st.ia[t];
      const int ia_hi = __double2hiint(ia), ia_lo = __double2loint(ia);
      double F0 = 0.0, F1 = 0.0, F2 = 0.0;
#pragma unroll
      for (int k = 0; k < RDYB_ELL_WIDTH; ++k) {
        if (k < width) {
          const uint32_t ce = st.ce[k][t];
          const int e = min((int)(ce & 0x7fffu), ME - 1);                  /* empty slot (0xffff) -> the zero flux entry */
          const int flip = (int)((~ce & 0x8000u) << 16);                   /* left cell: -1/area, right cell: +1/area */
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
    }
    if (EULER && q.push_src && tile < q.wait_tile_end) { /* fused halo push of a frontier tile's border cells (uniform per CTA; frontier tiles come first) */
      const int tn = p.tile_order ? __ldg(p.tile_order + tile) : tile;
      const int2 sd = __ldg(q.push_range + tn);
      if (sd.y > 0) {
        fixed_consumer_sync<NCONS>(); /* every cell of the tile has been written */
        for (int k = tid; k < sd.y; k += NCONS) {
          const int c = q.push_src[sd.x + k];
          double *d = q.push_dst[sd.x + k];
          const int64_t rs = q.push_rs[sd.x + k];
          d[0] = __ldcg(a.o0 + c); d[rs] = __ldcg(a.o1 + c); d[2 * rs] = __ldcg(a.o2 + c);
        }
        __threadfence_system(); /* the peer stores above before the counters below */
        fixed_consumer_sync<NCONS>();
        if (tid == 0) {
          const unsigned int done = atomicAdd(q.push_counter, 1u) + 1u;
          if (done == q.push_target) { /* the last tile with sends of this launch: every other CTA fenced before its atomicAdd */
            __threadfence_system();
            for (int pp = 0; pp < q.push_peers; ++pp) *reinterpret_cast<volatile unsigned int *>(q.push_flag[pp]) = q.push_epoch;
            __threadfence_system();
          }
        }
      }
    }
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
    fixed_consumer_sync<NCONS>();
    if (tid == 0) {
      for (int w = 1; w < NCONS / 32; ++w)
        if (S.wval[w] > cmax || (S.wval[w] == cmax && S.wargs[w] >= 0 && (carg < 0 || S.wargs[w] < carg))) { cmax = S.wval[w]; carg = S.wargs[w]; }
      const int slot = q.tile_begin + blockIdx.x; /* tile numbers of different launches do not overlap, so neither do the slots */
      const long long arg = carg;
      if (cmax > a.cmax[slot] || (cmax == a.cmax[slot] && arg >= 0 && (a.carg[slot] < 0 || arg < a.carg[slot]))) { a.cmax[slot] = cmax; a.carg[slot] = arg; }
    }
  }
}

/* ---------------- host side ---------------- */
template <int TC, int HM, int ME, int NCONS, int NS, int MINB, bool HR, int SRC, bool EULER, bool COURANT, int PM>
static int fixed_launch_t(const FixedArgs &q, int ntiles, cudaStream_t stream) {
  static PersistentCache cache = {};
  auto kern = swe_tiled_fixed_kernel<TC, HM, ME, NCONS, NS, MINB, HR, SRC, EULER, COURANT, PM>;
  const size_t smem = sizeof(FixedSmem<TC, HM, ME, NS, HR, PM>);
  int grid = 0;
  RDYB_CALL(persistent_grid(kern, NCONS + 32, smem, ntiles, cache, "fixed tile kernel", &grid));
  kern<<<grid, NCONS + 32, smem, stream>>>(q);
  return RDYB_SUCCESS;
}

/* which variant computes the primitive state (see FixedSmem): RDYB_K4_PRIM = 0 | 1 (A/B runs; RDYB_K4_PRIM_PASS=1 is the old spelling of 0) */
static inline int fixed_prim_mode() {
  static int v = -2;
  if (v == -2) {
    const char *e = getenv("RDYB_K4_PRIM"), *o = getenv("RDYB_K4_PRIM_PASS");
    v = 1;
    if (e && (e[0] == '0' || e[0] == '1')) v = e[0] - '0';
    else if (o && o[0] == '1') v = 0;
  }
  return v;
}
template <int TC, int HM, int ME, int NCONS, int NS, int MINB, bool HR, int PM>
static int fixed_launch_cfg2(const FixedArgs &q, int nt, bool xq, bool euler, bool courant, cudaStream_t stream) {
#define FL(SRCv)                                                                                                                   \
  (euler ? (courant ? fixed_launch_t<TC, HM, ME, NCONS, NS, MINB, HR, SRCv, true, true, PM>(q, nt, stream)                              \
                    : fixed_launch_t<TC, HM, ME, NCONS, NS, MINB, HR, SRCv, true, false, PM>(q, nt, stream))                            \
         : (courant ? fixed_launch_t<TC, HM, ME, NCONS, NS, MINB, HR, SRCv, false, true, PM>(q, nt, stream)                             \
                    : fixed_launch_t<TC, HM, ME, NCONS, NS, MINB, HR, SRCv, false, false, PM>(q, nt, stream)))
  return xq ? FL(1) : FL(0);
#undef FL
}
template <int TC, int HM, int ME, int NCONS, int NS, int MINB, bool HR>
static int fixed_launch_cfg(const FixedArgs &q, int nt, bool xq, bool euler, bool courant, cudaStream_t stream) {
  return fixed_prim_mode() == 0 ? fixed_launch_cfg2<TC, HM, ME, NCONS, NS, MINB, HR, 0>(q, nt, xq, euler, courant, stream)
                                : fixed_launch_cfg2<TC, HM, ME, NCONS, NS, MINB, HR, 1>(q, nt, xq, euler, courant, stream);
}

/* the compiled tile shapes: {tile cells, halo capacity, edge capacity (the last entry is the zero flux), consumer threads,
 * pipeline stages, CTAs/SM} */
#define RDYB_FIXED_A 256, 96, 576, 256, 2, 2 /* quads and mixed meshes (2 edges per cell) */
#define RDYB_FIXED_T 256, 80, 448, 256, 3, 2 /* triangle meshes (1.6 edges per cell): room for a third stage */
#define RDYB_FIXED_S 224, 80, 400, 224, 2, 3 /* triangle meshes, 7 consumer warps: three CTAs (21 consumer warps) per SM */

static inline bool fixed_fits_t(int tc, int mh, int me) { return tc == 256 && mh <= 80 && me <= 446; }
static inline bool fixed_fits_a(int tc, int mh, int me) { return tc == 256 && mh <= 96 && me <= 574; }
static inline bool fixed_fits_s(int tc, int mh, int me) { return tc == 224 && mh <= 80 && me <= 398; }
/* true when K4 has a compiled shape for tiles of tc cells with at most mh halo cells and me edges, and for this option set */
static inline bool tiled_fixed_applicable(int tc, int mh, int me, const SweArgs &a) {
  if (a.s1 || a.s2) return false; /* momentum sources: K3 */
  return fixed_fits_t(tc, mh, me) || fixed_fits_a(tc, mh, me) || fixed_fits_s(tc, mh, me);
}

struct FixedWait {
  int32_t peers;
  unsigned int epoch;
  const unsigned int *flags;
  int *status;
  long long timeout = 40000000000ll;
  unsigned long long *wait_cycles = nullptr;
  /* fused push (optional): see FixedArgs */
  const int32_t *order = nullptr; /* launch order with the frontier tiles first */
  const int2 *push_range = nullptr;
  const int32_t *push_src = nullptr;
  double *const *push_dst = nullptr;
  const int32_t *push_rs = nullptr;
  unsigned int *push_counter = nullptr;
  unsigned int push_target = 0, push_epoch = 0;
  unsigned int *const *push_flag = nullptr;
};

static inline int tiled_fixed_launch(const TiledDev &d, const RdybTiledLayout &T, const RdybConfig &cfg, const SweArgs &a, int32_t n_interior,
                                     bool euler, bool courant, int max_halo, int max_edges, cudaStream_t stream, const FixedWait *wait = nullptr) {
  FixedArgs q;
  TiledArgs &p = q.t;
  p.a = a;
  p.tile_desc = d.tile_desc; p.tile_cell_begin = d.tile_cell_begin; p.halo_offsets = d.halo_offsets; p.halo_cells = d.halo_cells;
  p.edge_offsets = d.edge_offsets; p.edge_bidx = d.edge_bidx; p.edge_lr = d.edge_lr; p.edge_sn = d.edge_sn; p.edge_cn = d.edge_cn;
  p.edge_len = d.edge_len; p.cell_edges = d.cell_edges;
  p.tile0 = 0; p.max_cells = T.tile_cells; p.max_staged = 0; p.max_edges = 0;
  { const char *dbg = getenv("RDYB_DEBUG_COPYONLY"); if (dbg && dbg[0] == '1') p.tile0 = 1; }
  q.tile_begin = a.begin < n_interior ? 0 : T.num_interior_tiles;
  q.tile_end = (a.end <= n_interior && a.begin < n_interior) ? T.num_interior_tiles : T.num_tiles;
  p.tile_order = d.order;
  if (d.order) { if (d.sub_begin > q.tile_begin) q.tile_begin = d.sub_begin; if (d.sub_end < q.tile_end) q.tile_end = d.sub_end; }
  if (q.tile_end <= q.tile_begin) return RDYB_SUCCESS;
  if (wait && wait->peers > 0 && !d.order && q.tile_end > T.num_interior_tiles) {
    q.wait_tile = T.num_interior_tiles; q.wait_tile_end = T.num_tiles;
    q.wait_peers = wait->peers; q.wait_epoch = wait->epoch; q.wait_flags = wait->flags; q.wait_status = wait->status;
    q.wait_timeout = wait->timeout; q.wait_cycles = wait->wait_cycles;
    if (wait->push_src && euler && q.tile_begin == 0 && q.tile_end == T.num_tiles) { /* whole range, frontier tiles first */
      p.tile_order = wait->order;
      q.wait_tile = 0; q.wait_tile_end = T.num_tiles - T.num_interior_tiles;
      q.push_range = wait->push_range; q.push_src = wait->push_src; q.push_dst = wait->push_dst; q.push_rs = wait->push_rs; q.push_counter = wait->push_counter;
      q.push_target = wait->push_target; q.push_epoch = wait->push_epoch; q.push_peers = wait->peers;
      for (int pp = 0; pp < wait->peers && pp < RDYB_MAX_PEERS; ++pp) q.push_flag[pp] = wait->push_flag[pp];
    }
  }
  const bool hr = cfg.well_balancing == RDYB_WELL_BALANCING_HR;
  const bool xq = cfg.source_method == RDYB_SOURCE_IMPLICIT_XQ2018;
  const int nt = q.tile_end - q.tile_begin;
  const char *env = getenv("RDYB_FIXED_SHAPE"); /* experiments: "A" forces the 2-stage shape */
  if (fixed_fits_s(T.tile_cells, max_halo, max_edges)) return hr ? fixed_launch_cfg<RDYB_FIXED_S, true>(q, nt, xq, euler, courant, stream) : fixed_launch_cfg<RDYB_FIXED_S, false>(q, nt, xq, euler, courant, stream);
  const bool use_t = fixed_fits_t(T.tile_cells, max_halo, max_edges) && !(env && env[0] == 'A');
  if (use_t) return hr ? fixed_launch_cfg<RDYB_FIXED_T, true>(q, nt, xq, euler, courant, stream) : fixed_launch_cfg<RDYB_FIXED_T, false>(q, nt, xq, euler, courant, stream);
  return hr ? fixed_launch_cfg<RDYB_FIXED_A, true>(q, nt, xq, euler, courant, stream) : fixed_launch_cfg<RDYB_FIXED_A, false>(q, nt, xq, euler, courant, stream);
}
#endif
