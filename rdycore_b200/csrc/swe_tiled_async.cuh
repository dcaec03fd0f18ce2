/* swe_tiled_async.cuh — K3: persistent, warp-specialised, bulk-async (TMA) staged tile kernel (sm_100a).
 *
 * Same three phases per tile as K2 (swe_tiled.cuh) but the loads are taken off the compute warps:
 *   - the grid is persistent (<= 2 CTAs per SM), each CTA walks tiles blockIdx.x, blockIdx.x + gridDim.x, ...;
 *   - one PRODUCER warp per CTA stages tile n+1 into a second shared-memory buffer while the 8 CONSUMER warps compute
 *     tile n: every per-tile array slice (state, 1/area, slopes, Manning n, rain, per-cell edge slots, per-edge
 *     endpoints / normal / length) is contiguous in the device layout and is fetched with one `cp.async.bulk`
 *     (TMA 1-D bulk copy, SASS UBLKCP) that signals an mbarrier with its byte count; the tile's halo cells are gathered
 *     with 8-byte `cp.async` (LDGSTS) tracked by the same mbarrier;
 *   - consumers wait on the "full" mbarrier, compute from shared memory only, write u + dt*F (or F) coalesced, and
 *     release the buffer through the "empty" mbarrier.
 * No float atomics; shared edges are evaluated with identical inputs and orientation in both tiles (exact conservation).
 */
#ifndef RDYB_SWE_TILED_ASYNC_CUH
#define RDYB_SWE_TILED_ASYNC_CUH
#include "swe_tiled.cuh"

#define RDYB_ASYNC_CONSUMERS 256
#define RDYB_ASYNC_THREADS (RDYB_ASYNC_CONSUMERS + 32)

struct AsyncArgs {
  TiledArgs t;
  int32_t tile_begin, tile_end; /* tiles of this launch */
  int32_t TC;                   /* tile capacity = local index of the first halo cell */
  int32_t CS;                   /* staged-cell capacity (TC + max halo, even) */
  int32_t ME;                   /* edge capacity (multiple of 4) */
  int32_t stage_bytes, scratch_off; /* byte size of one stage, byte offset of the scratch area */
};

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
/* try_wait suspends the thread in hardware until the phase completes or a time limit passes; without a hint that limit is short and
 * a waiting warp keeps re-issuing SYNCS + BRA + YIELD - 11 % (Roe) to 18 % (HR) of all issued instructions of K4 in the ncu capture of
 * round 2 (profiles/r2_ncu_full_swe_tiled_fixed_*.txt).  The hint (ns) lengthens the suspension; completion still wakes the thread
 * at once.  Measured: no change of the step time (0.3041 vs 0.3050 ms Roe, 0.3661 vs 0.3637 ms HR) - the re-issued instructions
 * fill slots nobody else wanted.  Kept: it cannot hurt and spares the scheduler. */
#ifndef RDYB_MBAR_SUSPEND_NS
#define RDYB_MBAR_SUSPEND_NS 20000
#endif
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, %2;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity), "r"((uint32_t)RDYB_MBAR_SUSPEND_NS)
      : "memory");
}
/* TMA 1-D bulk copy global -> shared, completion counted in bytes on `bar` */
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes, uint64_t *bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes),
               "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void cp_async8(void *dst, const void *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_arrive_noinc(uint64_t *bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void consumer_sync() { asm volatile("bar.sync 1, %0;" ::"n"(RDYB_ASYNC_CONSUMERS) : "memory"); }

/* pointers into one stage buffer */
struct StageView {
  double *h, *hu, *hv, *ia, *z;    /* [CS] */
  double *dzdx, *dzdy, *mann, *s0, *s1, *s2; /* [TC] */
  double *esn, *ecn, *elen;        /* [ME] */
  uint32_t *elr;                   /* [ME] */
  uint16_t *ce;                    /* [4][TC] */
  int32_t *desc;                   /* [8] */
};

template <bool HR>
__device__ __forceinline__ StageView stage_view(char *base, int TC, int CS, int ME) {
  StageView v;
  double *d = reinterpret_cast<double *>(base);
  v.h = d; v.hu = v.h + CS; v.hv = v.hu + CS; v.ia = v.hv + CS; v.z = v.ia + CS;
  d = v.z + (HR ? CS : 0);
  v.dzdx = d; v.dzdy = v.dzdx + TC; v.mann = v.dzdy + TC; v.s0 = v.mann + TC; v.s1 = v.s0 + TC; v.s2 = v.s1 + TC;
  d = v.s2 + TC;
  v.esn = d; v.ecn = v.esn + ME; v.elen = v.ecn + ME;
  v.elr = reinterpret_cast<uint32_t *>(v.elen + ME);
  v.ce = reinterpret_cast<uint16_t *>(v.elr + ME);
  v.desc = reinterpret_cast<int32_t *>(v.ce + 4 * TC);
  return v;
}
template <bool HR>
__host__ __device__ inline size_t stage_bytes_for(int TC, int CS, int ME) {
  return sizeof(double) * ((size_t)CS * (HR ? 5 : 4) + 6 * (size_t)TC + 3 * (size_t)ME) + 4 * (size_t)ME + 2 * 4 * (size_t)TC + 32;
}

template <bool HR, int SRC, bool EULER, bool COURANT>
__global__ void __launch_bounds__(RDYB_ASYNC_THREADS, 2) swe_tiled_async_kernel(const AsyncArgs q) {
  extern __shared__ __align__(128) char smem_raw[];
  __shared__ uint64_t full_bar[2], empty_bar[2];
  const TiledArgs &p = q.t;
  const SweArgs &a = p.a;
  const int TC = q.TC, CS = q.CS, ME = q.ME;
  const bool mom = a.s1 != nullptr;
  if (threadIdx.x == 0) {
    mbar_init(&full_bar[0], 33); mbar_init(&full_bar[1], 33);
    mbar_init(&empty_bar[0], RDYB_ASYNC_CONSUMERS); mbar_init(&empty_bar[1], RDYB_ASYNC_CONSUMERS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == RDYB_ASYNC_CONSUMERS / 32) {
    /* ================= producer warp ================= */
    int it = 0;
    for (int tile = q.tile_begin + blockIdx.x; tile < q.tile_end; tile += gridDim.x, ++it) {
      const int s = it & 1;
      mbar_wait(&empty_bar[s], ((it >> 1) & 1) ^ 1);
      const StageView v = stage_view<HR>(smem_raw + (size_t)s * q.stage_bytes, TC, CS, ME);
      const int tn = p.tile_order ? __ldg(p.tile_order + tile) : tile; /* launch position -> tile */
      const int4 d0 = *reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)tn);
      const int2 d1 = *reinterpret_cast<const int2 *>(p.tile_desc + 8 * (size_t)tn + 4);
      const int cb = d0.x, ncell = d0.y, eo = d0.z, nedge = d0.w, ho = d1.x, nhalo = d1.y;
      if (lane == 0) {
        v.desc[0] = cb; v.desc[1] = ncell; v.desc[2] = eo; v.desc[3] = nedge; v.desc[4] = ho; v.desc[5] = nhalo;
        const uint32_t bc = (uint32_t)((ncell + 1) & ~1) * 8u;  /* per-cell double slices */
        const uint32_t bh = (uint32_t)((ncell + 7) & ~7) * 2u;  /* per-cell uint16 slices */
        const uint32_t be = (uint32_t)((nedge + 1) & ~1) * 8u;  /* per-edge double slices */
        const uint32_t bl = (uint32_t)((nedge + 3) & ~3) * 4u;  /* per-edge uint32 slice */
        const int ncd = 3 + 1 + (HR ? 1 : 2) + 1 + 1 + (mom ? 2 : 0);
        mbar_expect_tx(&full_bar[s], bc * ncd + bh * a.width + be * 3 + bl);
        bulk_g2s(v.h, a.h + cb, bc, &full_bar[s]);
        bulk_g2s(v.hu, a.hu + cb, bc, &full_bar[s]);
        bulk_g2s(v.hv, a.hv + cb, bc, &full_bar[s]);
        bulk_g2s(v.ia, a.inv_area + cb, bc, &full_bar[s]);
        if (HR) bulk_g2s(v.z, a.zc + cb, bc, &full_bar[s]);
        else { bulk_g2s(v.dzdx, a.dzdx + cb, bc, &full_bar[s]); bulk_g2s(v.dzdy, a.dzdy + cb, bc, &full_bar[s]); }
        bulk_g2s(v.mann, a.mann + cb, bc, &full_bar[s]);
        bulk_g2s(v.s0, a.s0 + cb, bc, &full_bar[s]);
        if (mom) { bulk_g2s(v.s1, a.s1 + cb, bc, &full_bar[s]); bulk_g2s(v.s2, a.s2 + cb, bc, &full_bar[s]); }
        for (int k = 0; k < a.width; ++k) bulk_g2s(v.ce + (size_t)k * TC, p.cell_edges + (size_t)k * a.stride + cb, bh, &full_bar[s]);
        bulk_g2s(v.esn, p.edge_sn + eo, be, &full_bar[s]);
        bulk_g2s(v.ecn, p.edge_cn + eo, be, &full_bar[s]);
        bulk_g2s(v.elen, p.edge_len + eo, be, &full_bar[s]);
        bulk_g2s(v.elr, p.edge_lr + eo, bl, &full_bar[s]);
      }
      /* halo gather: 8-byte async copies straight into the staged arrays at local index TC + t */
      for (int t = lane; t < nhalo; t += 32) {
        const int c = p.halo_cells[ho + t];
        cp_async8(v.h + TC + t, a.h + c);
        cp_async8(v.hu + TC + t, a.hu + c);
        cp_async8(v.hv + TC + t, a.hv + c);
        cp_async8(v.ia + TC + t, a.inv_area + c);
        if (HR) cp_async8(v.z + TC + t, a.zc + c);
      }
      cp_async_arrive_noinc(&full_bar[s]);
    }
    return;
  }

  /* ================= consumer warps ================= */
  char *scratch = smem_raw + q.scratch_off;
  double *su = reinterpret_cast<double *>(scratch), *sv = su + CS, *ssq = sv + CS;
  double *sf0 = ssq + CS, *sf1 = sf0 + ME, *sf2 = sf1 + ME;
  double *c1l = sf2 + ME, *c2l = c1l + ME, *c1r = c2l + ME, *c2r = c1r + ME; /* HR only */
  const double g = RDYB_GRAVITY, tiny = a.tiny_h;
  const int tid = threadIdx.x;
  double cmax = 0.0;
  long long carg = -1;
  int it = 0;
  for (int tile = q.tile_begin + blockIdx.x; tile < q.tile_end; tile += gridDim.x, ++it) {
    const int s = it & 1;
    mbar_wait(&full_bar[s], (it >> 1) & 1);
    const StageView v = stage_view<HR>(smem_raw + (size_t)s * q.stage_bytes, TC, CS, ME);
    const int cb = v.desc[0], ncell = v.desc[1], eo = v.desc[2], nedge = v.desc[3], nhalo = v.desc[5];

    /* ---- phase 1: primitive state of tile cells [0,ncell) and halo cells [TC,TC+nhalo) ---- */
    for (int t = tid; t < ncell + nhalo; t += RDYB_ASYNC_CONSUMERS) {
      const int x = t < ncell ? t : TC + (t - ncell);
      const double h = v.h[x];
      double u, w;
      swe_velocities<HR>(h, v.hu[x], v.hv[x], tiny, a.ha2, u, w);
      su[x] = u; sv[x] = w;
      if (!HR) ssq[x] = swe_sqrt(h);
    }
    consumer_sync();

    /* ---- phase 2: one flux per tile edge ---- */
    for (int e = tid; e < nedge; e += RDYB_ASYNC_CONSUMERS) {
      const uint32_t lr = v.elr[e];
      const int l = lr & 0xffffu, r = lr >> 16;
      const double sn = v.esn[e], cn = v.ecn[e], len = v.elen[e];
      double hl = v.h[l], ul = su[l], vl = sv[l];
      RoeFlux f = {0.0, 0.0, 0.0, 0.0};
      double k1l = 0.0, k2l = 0.0, k1r = 0.0, k2r = 0.0;
      if (r != 0xffff) {
        double hr = v.h[r];
        if (!(hl < tiny && hr < tiny)) {
          const double ur = su[r], vr = sv[r];
          bool do_flux = true;
          if (HR) {
            const double zl = v.z[l], zr = v.z[r], zmax = fmax(zl, zr);
            const double hls = fmax(0.0, (hl + zl) - zmax), hrs = fmax(0.0, (hr + zr) - zmax);
            const double corl = 0.5 * g * (hl * hl - hls * hls) * len, corr = 0.5 * g * (hr * hr - hrs * hrs) * len;
            k1l = corl * cn; k2l = corl * sn; k1r = corr * cn; k2r = corr * sn;
            hl = hls; hr = hrs;
            do_flux = (hls > tiny || hrs > tiny);
          }
          if (do_flux) {
            f = HR ? swe_roe_flux(hl, ul, vl, swe_sqrt(hl), hr, ur, vr, swe_sqrt(hr), sn, cn)
                   : swe_roe_flux(hl, ul, vl, ssq[l], hr, ur, vr, ssq[r], sn, cn);
            if (COURANT) {
              const double cnum = f.amax * len * fmax(v.ia[l], v.ia[r]) * a.dt;
              if (cnum > cmax) { cmax = cnum; carg = (long long)eo + e; }
            }
          }
        }
      } else {
        const int j = p.edge_bidx[eo + e];
        const double h = hl;
        swe_velocities<false>(h, v.hu[l], v.hv[l], tiny, a.ha2, ul, vl);
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
          if (COURANT) {
            const double cnum = f.amax * len * v.ia[l] * a.dt;
            if (cnum > cmax) { cmax = cnum; carg = (long long)eo + e; }
          }
        }
        if (a.bflux) {
          a.bflux[j] = f.f0; a.bflux[a.nbe + j] = f.f1; a.bflux[2 * a.nbe + j] = f.f2;
          a.bacc[j] += a.dt * f.f0; a.bacc[a.nbe + j] += a.dt * f.f1; a.bacc[2 * a.nbe + j] += a.dt * f.f2;
        }
      }
      sf0[e] = f.f0 * len; sf1[e] = f.f1 * len; sf2[e] = f.f2 * len;
      if (HR) { c1l[e] = k1l; c2l[e] = k2l; c1r[e] = k1r; c2r[e] = k2r; }
    }
    consumer_sync();

    /* ---- phase 3: per-cell flux sum, sources, update ---- */
    for (int t = tid; t < ncell; t += RDYB_ASYNC_CONSUMERS) {
      const int i = cb + t;
      const double h = v.h[t], hu = v.hu[t], hv = v.hv[t], ia = v.ia[t];
      double F0 = 0.0, F1 = 0.0, F2 = 0.0;
#pragma unroll
      for (int k = 0; k < RDYB_ELL_WIDTH; ++k) {
        if (k >= a.width) break;
        const uint32_t ce = v.ce[k * TC + t];
        if (ce == 0xffffu) continue;
        const int e = ce & 0x7fffu;
        const bool right = (ce & 0x8000u) != 0;
        const double sgn = right ? ia : -ia;
        F0 += sf0[e] * sgn; F1 += sf1[e] * sgn; F2 += sf2[e] * sgn;
        if (HR) { F1 += (right ? c1r[e] : c1l[e]) * sgn; F2 += (right ? c2r[e] : c2l[e]) * sgn; }
      }
      if (a.fdiv) { a.fdiv[i] = F0; a.fdiv[a.stride + i] = F1; a.fdiv[2 * a.stride + i] = F2; }
      double bedx = 0.0, bedy = 0.0, tbx = 0.0, tby = 0.0;
      if (!HR) { bedx = v.dzdx[t] * g * h; bedy = v.dzdy[t] * g * h; }
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

  if (COURANT && a.cmax) {
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
      const int slot = q.tile_begin + blockIdx.x; /* tile numbers of different launches do not overlap, so neither do the slots */
      if (cmax > a.cmax[slot] || (cmax == a.cmax[slot] && carg >= 0 && (a.carg[slot] < 0 || carg < a.carg[slot]))) { a.cmax[slot] = cmax; a.carg[slot] = carg; }
    }
  }
}

/* ---------------- host side ---------------- */
template <bool HR, int SRC, bool EULER, bool COURANT>
static int async_launch_t(AsyncArgs &q, int ntiles, size_t smem, cudaStream_t stream) {
  static PersistentCache cache = {};
  auto kern = swe_tiled_async_kernel<HR, SRC, EULER, COURANT>;
  int grid = 0;
  RDYB_CALL(persistent_grid(kern, RDYB_ASYNC_THREADS, smem, ntiles, cache, "async tile kernel", &grid));
  kern<<<grid, RDYB_ASYNC_THREADS, smem, stream>>>(q);
  return RDYB_SUCCESS;
}

static inline int tiled_async_launch(const TiledDev &d, const RdybTiledLayout &T, const RdybConfig &cfg, const SweArgs &a, int32_t n_interior,
                                     bool euler, bool courant, cudaStream_t stream) {
  AsyncArgs q;
  TiledArgs &p = q.t;
  p.a = a;
  p.tile_desc = d.tile_desc; p.tile_cell_begin = d.tile_cell_begin; p.halo_offsets = d.halo_offsets; p.halo_cells = d.halo_cells;
  p.edge_offsets = d.edge_offsets; p.edge_bidx = d.edge_bidx; p.edge_lr = d.edge_lr; p.edge_sn = d.edge_sn; p.edge_cn = d.edge_cn;
  p.edge_len = d.edge_len; p.cell_edges = d.cell_edges;
  p.tile0 = 0; p.max_cells = T.tile_cells; p.max_staged = 0; p.max_edges = 0;
  q.tile_begin = a.begin < n_interior ? 0 : T.num_interior_tiles;
  q.tile_end = (a.end <= n_interior && a.begin < n_interior) ? T.num_interior_tiles : T.num_tiles;
  p.tile_order = d.order;
  if (d.order) { if (d.sub_begin > q.tile_begin) q.tile_begin = d.sub_begin; if (d.sub_end < q.tile_end) q.tile_end = d.sub_end; }
  if (q.tile_end <= q.tile_begin) return RDYB_SUCCESS;
  const bool hr = cfg.well_balancing == RDYB_WELL_BALANCING_HR;
  const bool xq = cfg.source_method == RDYB_SOURCE_IMPLICIT_XQ2018;
  q.TC = T.tile_cells;
  q.CS = T.tile_cells + ((T.max_halo + 1) & ~1);
  q.ME = (T.max_edges + 3) & ~3;
  const size_t sb = hr ? stage_bytes_for<true>(q.TC, q.CS, q.ME) : stage_bytes_for<false>(q.TC, q.CS, q.ME);
  q.stage_bytes = (int32_t)((sb + 127) & ~(size_t)127);
  q.scratch_off = 2 * q.stage_bytes;
  const size_t smem = (size_t)q.scratch_off + sizeof(double) * (3 * (size_t)q.CS + (size_t)q.ME * (hr ? 7 : 3));
  RDYB_CHECK(smem <= 227 * 1024, RDYB_ERR_SUP, "async tile kernel: tile needs %zu bytes of shared memory (> 227 KB); use smaller tiles", smem);
  const int nt = q.tile_end - q.tile_begin;
#define AL(HRv, SRCv)                                                                                                     \
  (euler ? (courant ? async_launch_t<HRv, SRCv, true, true>(q, nt, smem, stream) : async_launch_t<HRv, SRCv, true, false>(q, nt, smem, stream)) \
         : (courant ? async_launch_t<HRv, SRCv, false, true>(q, nt, smem, stream) : async_launch_t<HRv, SRCv, false, false>(q, nt, smem, stream)))
  int ierr = hr ? (xq ? AL(true, 1) : AL(true, 0)) : (xq ? AL(false, 1) : AL(false, 0));
#undef AL
  return ierr;
}
#endif
