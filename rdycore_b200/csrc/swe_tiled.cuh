/* swe_tiled.cuh — K2 "tiled" kernel of the SWE right-hand side + explicit stage update (sm_100a).
 *
 * One thread block per tile of contiguous (Hilbert-ordered) cells, three phases separated by __syncthreads():
 *   1. stage: the tile's cells (coalesced) and its halo cells (gathered through L2) are loaded once, their primitive
 *      state (h,u,v) and 1/area are computed once per cell and kept in shared memory;
 *   2. edges: one thread per tile edge evaluates the Roe (or hydrostatic-reconstruction) flux ONCE from shared memory,
 *      applies boundary conditions on boundary edges, tracks the Courant number, and stores flux*length in shared memory;
 *   3. cells: one thread per cell sums its <= 4 edge fluxes in the reference's accumulation order (deterministic, no
 *      atomics), adds bed slope / friction / external sources and writes F or u + dt*F (coalesced).
 * Edges shared by two tiles are evaluated by both with the same left/right orientation and the same inputs, so both get
 * the same bits and mass is conserved exactly.  Same reference functions as K1 (see swe_kernels.cuh).
 */
#ifndef RDYB_SWE_TILED_CUH
#define RDYB_SWE_TILED_CUH
#include <cuda_runtime.h>

#include "device_layout.h"
#include "rdyb200.h"
#include "rdyb_error.h"
#include "swe_kernels.cuh"

#define RDYB_TILE_THREADS 256

struct TiledDev {
  int32_t *tile_cell_begin, *tile_desc, *halo_offsets, *halo_cells, *edge_offsets, *edge_bidx;
  uint32_t *edge_lr;
  double *edge_sn, *edge_cn, *edge_len;
  uint16_t *cell_edges;
  /* optional launch order (pipelined host-Vec apply): positions [sub_begin, sub_end) of `order` name the tiles to run */
  const int32_t *order;
  int32_t sub_begin, sub_end;
};

struct TiledArgs {
  SweArgs a;
  const int32_t *__restrict__ tile_desc;
  const int32_t *__restrict__ tile_cell_begin, *__restrict__ halo_offsets, *__restrict__ halo_cells, *__restrict__ edge_offsets;
  const int32_t *__restrict__ edge_bidx;
  const uint32_t *__restrict__ edge_lr;
  const double *__restrict__ edge_sn, *__restrict__ edge_cn, *__restrict__ edge_len;
  const uint16_t *__restrict__ cell_edges;
  const int32_t *__restrict__ tile_order; /* NULL: tiles run in index order; else launch position -> tile (K3, K4) */
  int32_t tile0;      /* first tile of this launch */
  int32_t max_cells;  /* shared-memory capacities */
  int32_t max_staged; /* tile cells + halo cells */
  int32_t max_edges;
};

template <bool HR, int SRC, bool EULER, bool COURANT>
__global__ void __launch_bounds__(RDYB_TILE_THREADS, 2) swe_tiled_kernel(const TiledArgs p) {
  extern __shared__ double smem[];
  const SweArgs &a = p.a;
  const int tile = p.tile0 + blockIdx.x;
  const int4 d0 = *reinterpret_cast<const int4 *>(p.tile_desc + 8 * (size_t)tile);
  const int2 d1 = *reinterpret_cast<const int2 *>(p.tile_desc + 8 * (size_t)tile + 4);
  const int cb = d0.x, ncell = d0.y, eo = d0.z, nedge = d0.w, ho = d1.x, nhalo = d1.y;
  /* shared-memory carve-up */
  double *sh = smem, *su = sh + p.max_staged, *sv = su + p.max_staged, *sia = sv + p.max_staged, *ssq = sia + p.max_staged;
  double *sz = ssq + p.max_staged; /* HR only */
  double *shu = sz + (HR ? p.max_staged : 0), *shv = shu + p.max_cells;
  double *sf0 = shv + p.max_cells, *sf1 = sf0 + p.max_edges, *sf2 = sf1 + p.max_edges;
  double *c1l = sf2 + p.max_edges, *c2l = c1l + p.max_edges, *c1r = c2l + p.max_edges, *c2r = c1r + p.max_edges; /* HR only */
  const double g = RDYB_GRAVITY, tiny = a.tiny_h;

  /* ---- phase 1: stage tile cells and halo cells ---- */
  for (int t = threadIdx.x; t < ncell + nhalo; t += RDYB_TILE_THREADS) {
    const int c = t < ncell ? cb + t : p.halo_cells[ho + t - ncell];
    const int x = t < ncell ? t : p.max_cells + (t - ncell); /* halo cells sit at local index TC + k */
    const double h = a.h[c], hu = a.hu[c], hv = a.hv[c];
    double u, v;
    swe_velocities<HR>(h, hu, hv, tiny, a.ha2, u, v);
    sh[x] = h; su[x] = u; sv[x] = v; sia[x] = a.inv_area[c];
    if (HR) sz[x] = a.zc[c];
    else ssq[x] = swe_sqrt(h);
    if (t < ncell) { shu[t] = hu; shv[t] = hv; }
  }
  __syncthreads();

  /* ---- phase 2: one flux per tile edge ---- */
  double cmax = 0.0;
  long long carg = -1;
  for (int e = threadIdx.x; e < nedge; e += RDYB_TILE_THREADS) {
    const uint32_t lr = p.edge_lr[eo + e];
    const int l = lr & 0xffffu, r = lr >> 16;
    const double sn = p.edge_sn[eo + e], cn = p.edge_cn[eo + e], len = p.edge_len[eo + e];
    double hl = sh[l], ul = su[l], vl = sv[l];
    RoeFlux f = {0.0, 0.0, 0.0, 0.0};
    double k1l = 0.0, k2l = 0.0, k1r = 0.0, k2r = 0.0;
    if (r != 0xffff) {
      double hr = sh[r];
      if (!(hl < tiny && hr < tiny)) {
        const double ur = su[r], vr = sv[r];
        bool do_flux = true;
        if (HR) {
          const double zl = sz[l], zr = sz[r], zmax = fmax(zl, zr);
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
            const double cnum = f.amax * len * fmax(sia[l], sia[r]) * a.dt;
            if (cnum > cmax) { cmax = cnum; carg = (long long)eo + e; }
          }
        }
      }
    } else {
      /* boundary edge (left cell is a tile cell); non-strict wet test as in ApplyBoundaryFlux */
      const int j = p.edge_bidx[eo + e];
      const double h = hl;
      swe_velocities<false>(h, shu[l], shv[l], tiny, a.ha2, ul, vl);
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
          const double cnum = f.amax * len * sia[l] * a.dt;
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
  __syncthreads();

  /* ---- phase 3: per-cell flux sum, sources, update ---- */
  for (int t = threadIdx.x; t < ncell; t += RDYB_TILE_THREADS) {
    const int i = cb + t;
    const double h = sh[t], hu = shu[t], hv = shv[t], ia = sia[t];
    double F0 = 0.0, F1 = 0.0, F2 = 0.0;
#pragma unroll
    for (int k = 0; k < RDYB_ELL_WIDTH; ++k) {
      if (k >= a.width) break;
      const uint32_t ce = p.cell_edges[(size_t)k * a.stride + i];
      if (ce == 0xffffu) continue;
      const int e = ce & 0x7fffu;
      const bool right = (ce & 0x8000u) != 0;
      const double s = right ? ia : -ia; /* +1/A for the right cell, -1/A for the left cell */
      F0 += sf0[e] * s;
      F1 += sf1[e] * s;
      F2 += sf2[e] * s;
      if (HR) { F1 += (right ? c1r[e] : c1l[e]) * s; F2 += (right ? c2r[e] : c2l[e]) * s; }
    }
    if (a.fdiv) { a.fdiv[i] = F0; a.fdiv[a.stride + i] = F1; a.fdiv[2 * a.stride + i] = F2; }
    double bedx = 0.0, bedy = 0.0, tbx = 0.0, tby = 0.0;
    if (!HR) { bedx = a.dzdx[i] * g * h; bedy = a.dzdy[i] * g * h; }
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

  if (COURANT && a.cmax) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
      const double ov = __shfl_down_sync(0xffffffffu, cmax, off);
      const long long oa = __shfl_down_sync(0xffffffffu, carg, off);
      if (ov > cmax || (ov == cmax && oa >= 0 && (carg < 0 || oa < carg))) { cmax = ov; carg = oa; }
    }
    __shared__ double wval[RDYB_TILE_THREADS / 32];
    __shared__ long long warg[RDYB_TILE_THREADS / 32];
    if ((threadIdx.x & 31) == 0) { wval[threadIdx.x >> 5] = cmax; warg[threadIdx.x >> 5] = carg; }
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int w = 1; w < RDYB_TILE_THREADS / 32; ++w)
        if (wval[w] > cmax || (wval[w] == cmax && warg[w] >= 0 && (carg < 0 || warg[w] < carg))) { cmax = wval[w]; carg = warg[w]; }
      if (cmax > a.cmax[tile]) { a.cmax[tile] = cmax; a.carg[tile] = carg; }
    }
  }
}

/* ---------------- host side ---------------- */
static inline void tiled_dev_free(TiledDev *d) {
  void *ptrs[] = {d->tile_desc, d->tile_cell_begin, d->halo_offsets, d->halo_cells, d->edge_offsets, d->edge_bidx, d->edge_lr, d->edge_sn, d->edge_cn, d->edge_len, d->cell_edges};
  for (void *q : ptrs)
    if (q) cudaFree(q);
  memset(d, 0, sizeof(*d));
}

template <typename T>
static int tiled_up(T **dst, const T *src, size_t n) {
  cudaError_t e = cudaMalloc((void **)dst, sizeof(T) * (n > 0 ? n : 1));
  if (e == cudaSuccess && n) e = cudaMemcpy(*dst, src, sizeof(T) * n, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) return rdyb_fail(RDYB_ERR_CUDA, "CUDA error %s while uploading the tiled layout (%s)", cudaGetErrorName(e), cudaGetErrorString(e));
  return RDYB_SUCCESS;
}

static inline int tiled_dev_upload(const RdybTiledLayout *T, const RdybLayout *L, TiledDev *d) {
  memset(d, 0, sizeof(*d));
  const size_t nt = (size_t)T->num_tiles, ne = (size_t)T->total_edges;
  RDYB_CALL(tiled_up(&d->tile_cell_begin, T->tile_cell_begin, nt + 1));
  RDYB_CALL(tiled_up(&d->tile_desc, T->tile_desc, 8 * nt));
  RDYB_CALL(tiled_up(&d->halo_offsets, T->halo_offsets, nt + 1));
  RDYB_CALL(tiled_up(&d->halo_cells, T->halo_cells, (size_t)T->total_halo));
  RDYB_CALL(tiled_up(&d->edge_offsets, T->edge_offsets, nt + 1));
  RDYB_CALL(tiled_up(&d->edge_bidx, T->edge_bidx, ne));
  RDYB_CALL(tiled_up(&d->edge_lr, T->edge_lr, ne));
  RDYB_CALL(tiled_up(&d->edge_sn, T->edge_sn, ne));
  RDYB_CALL(tiled_up(&d->edge_cn, T->edge_cn, ne));
  RDYB_CALL(tiled_up(&d->edge_len, T->edge_len, ne));
  RDYB_CALL(tiled_up(&d->cell_edges, T->cell_edges, (size_t)RDYB_ELL_WIDTH * (size_t)L->stride));
  return RDYB_SUCCESS;
}


/* ---------------- grid of the persistent kernels (K3, K4, K3M) ----------------
 * One CTA per resident slot: SMs x CTAs/SM, cached PER DEVICE (cudaFuncSetAttribute and the occupancy are per device; an
 * operator may be created on any device of the process).  Caps, for experiments and tests:
 *   RDYB_CTAS_PER_SM=n   fewer resident CTAs per SM;
 *   rdyb_operator_set_max_ctas(op, n) / RDYB_MAX_CTAS=n   at most n CTAs in total, so that on a small mesh every CTA walks many
 *   tiles through all pipeline stages and both mbarrier parities - the regime of the 10 M-cell bench (tests/test_gpu_parity.py). */
#define RDYB_MAX_DEVICES 64
struct PersistentCache {
  size_t smem[RDYB_MAX_DEVICES];
  int bps[RDYB_MAX_DEVICES], sms[RDYB_MAX_DEVICES];
};
static thread_local int tl_rdyb_max_ctas = 0; /* the launching operator's cap (launch_cells sets it before every launch) */

template <typename Kern>
static int persistent_grid(Kern kern, int threads, size_t smem, int ntiles, PersistentCache &c, const char *what, int *grid) {
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= RDYB_MAX_DEVICES) return rdyb_fail(RDYB_ERR_SUP, "%s: device %d out of the supported range", what, dev);
  if (c.smem[dev] != smem || c.bps[dev] < 1) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return rdyb_fail(RDYB_ERR_CUDA, "%s: %zu bytes of shared memory per CTA not available (%s)", what, smem, cudaGetErrorString(e));
    cudaDeviceGetAttribute(&c.sms[dev], cudaDevAttrMultiProcessorCount, dev);
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c.bps[dev], kern, threads, smem);
    if (e != cudaSuccess || c.bps[dev] < 1) { c.bps[dev] = 0; return rdyb_fail(RDYB_ERR_CUDA, "%s does not fit on an SM with %zu bytes of shared memory", what, smem); }
    c.smem[dev] = smem;
  }
  const char *env = getenv("RDYB_CTAS_PER_SM");
  int bps = c.bps[dev];
  if (env && atoi(env) > 0 && atoi(env) < bps) bps = atoi(env);
  int g = c.sms[dev] * bps;
  int cap = tl_rdyb_max_ctas;
  const char *envc = getenv("RDYB_MAX_CTAS");
  if (cap <= 0 && envc && atoi(envc) > 0) cap = atoi(envc);
  if (cap > 0 && g > cap) g = cap;
  if (g > ntiles) g = ntiles;
  *grid = g;
  return RDYB_SUCCESS;
}

template <bool HR, int SRC, bool EULER, bool COURANT>
static int tiled_launch_t(const TiledArgs &p, int ntiles, size_t smem, cudaStream_t stream) {
  static size_t configured_dev[RDYB_MAX_DEVICES] = {}; /* per instantiation and device */
  int dev_ = 0;
  cudaGetDevice(&dev_);
  size_t &configured = configured_dev[dev_ >= 0 && dev_ < RDYB_MAX_DEVICES ? dev_ : 0];
  if (smem > 48 * 1024 && smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(swe_tiled_kernel<HR, SRC, EULER, COURANT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return rdyb_fail(RDYB_ERR_CUDA, "tiled kernel: %zu bytes of shared memory per tile not available (%s)", smem, cudaGetErrorString(e));
    configured = smem;
  }
  swe_tiled_kernel<HR, SRC, EULER, COURANT><<<ntiles, RDYB_TILE_THREADS, smem, stream>>>(p);
  return RDYB_SUCCESS;
}

/* `a` carries begin/end (a whole interior or frontier range, or everything) and all per-cell arrays */
static inline int tiled_launch(const TiledDev &d, const RdybTiledLayout &T, const RdybConfig &cfg, const SweArgs &a, int32_t n_interior, bool euler,
                               bool courant, cudaStream_t stream) {
  TiledArgs p;
  p.a = a;
  p.tile_desc = d.tile_desc; p.tile_cell_begin = d.tile_cell_begin; p.halo_offsets = d.halo_offsets; p.halo_cells = d.halo_cells; p.edge_offsets = d.edge_offsets;
  p.edge_bidx = d.edge_bidx; p.edge_lr = d.edge_lr; p.edge_sn = d.edge_sn; p.edge_cn = d.edge_cn; p.edge_len = d.edge_len;
  p.cell_edges = d.cell_edges;
  const int32_t t0 = a.begin < n_interior ? 0 : T.num_interior_tiles;
  const int32_t t1 = a.end <= n_interior && a.begin < n_interior ? T.num_interior_tiles : T.num_tiles;
  if (t1 <= t0) return RDYB_SUCCESS;
  p.tile0 = t0;
  p.max_cells = T.tile_cells; p.max_staged = T.tile_cells + T.max_halo; p.max_edges = T.max_edges;
  const bool hr = cfg.well_balancing == RDYB_WELL_BALANCING_HR;
  const bool xq = cfg.source_method == RDYB_SOURCE_IMPLICIT_XQ2018;
  const size_t smem = sizeof(double) * ((size_t)p.max_staged * (hr ? 6 : 5) + 2 * (size_t)p.max_cells + (size_t)p.max_edges * (hr ? 7 : 3));
  RDYB_CHECK(smem <= 227 * 1024, RDYB_ERR_SUP, "tiled kernel: tile needs %zu bytes of shared memory (> 227 KB); use smaller tiles", smem);
  const int nt = t1 - t0;
#define TL(HRv, SRCv)                                                                                         \
  (euler ? (courant ? tiled_launch_t<HRv, SRCv, true, true>(p, nt, smem, stream) : tiled_launch_t<HRv, SRCv, true, false>(p, nt, smem, stream)) \
         : (courant ? tiled_launch_t<HRv, SRCv, false, true>(p, nt, smem, stream) : tiled_launch_t<HRv, SRCv, false, false>(p, nt, smem, stream)))
  int ierr = hr ? (xq ? TL(true, 1) : TL(true, 0)) : (xq ? TL(false, 1) : TL(false, 0));
#undef TL
  return ierr;
}
#endif
