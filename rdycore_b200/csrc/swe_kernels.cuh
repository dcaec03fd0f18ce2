/* swe_kernels.cuh — sm_100a kernels of the SWE right-hand side + explicit stage update.
 *
 * K1 "gather": one thread per owned cell; the cell walks its ELL edge list (device_layout.h), gathers the neighbour's
 * state, evaluates the Roe flux with the edge's own left/right orientation (so both cells of an edge compute the same
 * bits and mass is conserved exactly), applies boundary conditions for boundary slots, then the bed-slope / friction /
 * external sources of the cell, and either writes F (RHS mode: ApplyPetscOperator, src/operator.c:656-672) or the new
 * state u + dt*F (Euler mode: + TSEULER).  No float atomics: the result is deterministic.
 *
 * Replaces, fused: ApplyInteriorFlux src/swe/swe_petsc.c:215-316, ApplyInteriorFluxHR :1000-1161, ApplyBoundaryFlux
 * :506-630, ApplySourceSemiImplicit :704-804, ApplySourceImplicitXQ2018 :816-932, VecZeroEntries src/rdysetup.c:1130,
 * the two VecCopy of src/operator.c:663,669 and the TS stage update.
 */
#ifndef RDYB_SWE_KERNELS_CUH
#define RDYB_SWE_KERNELS_CUH

#include <stdint.h>

#include "device_layout.h"
#include "rdyb200.h"
#include "swe_math.cuh"

#define RDYB_BLOCK 128

/* Device-resident time control of a coupling interval (rdyb_operator_advance_device): the step kernels of the device-side loop
 * read their time step, their turn (which state buffer is current) and whether the interval is over from here, so that
 * RDyAdvance needs no host round trip - not for the adaptive-dt rule (src/rdyadvance.c:303-343), not for the step count. */
struct AdvCtl {
  double t, t_end, dt, dt_step, rest, courant;
  long long steps;
  int nfull, k, total, cur, active, courant_valid;
};

struct SweArgs {
  /* state in (device order, SoA, length n_local) */
  const double *__restrict__ h, *__restrict__ hu, *__restrict__ hv;
  /* out: F (RHS mode) or new state (Euler mode), SoA */
  double *__restrict__ o0, *__restrict__ o1, *__restrict__ o2;
  const double *__restrict__ inv_area, *__restrict__ dzdx, *__restrict__ dzdy, *__restrict__ zc, *__restrict__ mann;
  const double *__restrict__ s0, *__restrict__ s1, *__restrict__ s2; /* external sources; s1,s2 may be NULL (all zero) */
  const int32_t *__restrict__ nbr;
  const double *__restrict__ esn, *__restrict__ ecn, *__restrict__ eslen;
  /* boundary-edge table */
  const int32_t *__restrict__ btype;
  const double *__restrict__ bval; /* [3][nbe] */
  double *__restrict__ bflux, *__restrict__ bacc; /* [3][nbe] or NULL */
  int32_t nbe;
  /* optional outputs (NULL = off): flux divergence and primitive variables, SoA [3][stride] */
  double *__restrict__ fdiv, *__restrict__ prim;
  /* Courant diagnostics: one slot per block (running max since reset), NULL = off */
  double *__restrict__ cmax;
  long long *__restrict__ carg;
  int32_t stride, width, begin, end;
  double dt, tiny_h, ha2, xq_thr;
  /* second order (MUSCL; K2M reads grad only, the per-slot arrays are for K1 and the gradient kernel): cell gradients SoA [6][stride] = d/dx (h,hu,hv), d/dy (h,hu,hv) incl. ghost cells, and the
   * per-slot least-squares weights / face offsets of device_layout.h; all NULL for first order */
  const double *__restrict__ grad;
  const double *__restrict__ gcx, *__restrict__ gcy, *__restrict__ dxs, *__restrict__ dys, *__restrict__ dxn, *__restrict__ dyn;
  /* device-side loop (K1 and K4 only): when set, dt is ctl->dt_step and the launch is a no-op unless the interval is active and
   * ctl->cur == ctl_buf (the launch reads state buffer ctl_buf) */
  const AdvCtl *ctl = nullptr;
  int32_t ctl_buf = 0;
  /* Bradford-Sanders (2002) well balancing, libCEED semantics (K1 only; NULL = off): water-surface elevation at the vertices
   * (bs2002_eta_vertex_kernel), vertex bed elevations, and per ELL slot the two vertices of the edge (RDyEdges.vertex_ids order) */
  const double *__restrict__ bs_eta = nullptr, *__restrict__ bs_vz = nullptr;
  const int32_t *__restrict__ bs_vb = nullptr, *__restrict__ bs_ve = nullptr;
};

/* ComputeDhv + the well-balancing term of SWERiemannFlux_Roe (src/swe/swe_fluxes_ceed.h:15-19, src/swe/swe_roe_flux_ceed.h:119-121):
 * g dhv^2 / 24, to be added as (cn, sn) times it to the momentum components of the edge flux */
__device__ __forceinline__ double swe_bs2002_term(const SweArgs &a, size_t q) {
  const int vb = a.bs_vb[q], ve = a.bs_ve[q];
  const double hv_beg = fmax(a.bs_eta[vb] - a.bs_vz[vb], 0.0), hv_end = fmax(a.bs_eta[ve] - a.bs_vz[ve], 0.0);
  const double dhv = hv_end - hv_beg;
  return RDYB_GRAVITY * (dhv * dhv) / 24.0;
}

/* SWEEtaVertex (src/swe/swe_well_balancing.h:30-83) gathered per vertex: eta[v] = sum over the cells around v, in ascending cell id,
 * of eta_cell(h; z1 <= z2 <= z3) / n_cells(v) (CreateCeedEtaVerticesOperator, src/operator_fluxes_ceed.c:745-859).  One thread per
 * vertex; no atomics, so the sum has a fixed order. */
__global__ void bs2002_eta_vertex_kernel(const double *__restrict__ h, const double *__restrict__ cz /* [3][stride] sorted vertex z per cell */, int32_t stride,
                                         const int32_t *__restrict__ vc_off, const int32_t *__restrict__ vc_cell, int32_t nv, double tiny_h,
                                         double *__restrict__ eta) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= nv) return;
  const int k0 = vc_off[v], k1 = vc_off[v + 1];
  const double wt = 1.0 / (double)(k1 - k0);
  double sum = 0.0;
  for (int k = k0; k < k1; ++k) {
    const int c = vc_cell[k];
    const double hc = h[c], z1 = cz[c], z2 = cz[(size_t)stride + c], z3 = cz[2 * (size_t)stride + c];
    const double h3 = z3 - (z1 + z2 + z3) / 3.0;
    const double h2 = z2 < z3 ? (z2 - z1) * (z2 - z1) / (3.0 * (z3 - z1)) : h3;
    double eta_cell = 0.0;
    if (hc <= tiny_h) eta_cell = z1; /* dry bed */
    else if (hc >= h3) eta_cell = (z1 + z2 + z3) / 3.0 + hc; /* all vertices submerged */
    else if (hc > 0.0 && hc <= h2) eta_cell = z1 + cbrt(3.0 * hc * (z2 - z1) * (z3 - z1)); /* one vertex submerged */
    else if (hc > h2 && hc < h3) { /* two vertices submerged */
      const double b = z3 - 3.0 * z1;
      const double c2 = z1 * z2 + z1 * z1 - z3 * z2 - 3.0 * hc * (z3 - z1);
      eta_cell = 0.5 * (-b + sqrt(b * b - 4.0 * c2));
    }
    sum += eta_cell * wt;
  }
  eta[v] = sum;
}

/* slope limiters of the MUSCL reconstruction, src/operator_fluxes_ceed.c:1100-1132: the extrapolated increment is limited
 * against half the cell-to-cell jump */
template <int RECON>
__device__ __forceinline__ double swe_limit_slope(double extrap, double half_dq) {
  if (RECON == RDYB_RECON_MUSCL_NONE) return extrap;
  if (extrap * half_dq <= 0.0) return 0.0;
  if (RECON == RDYB_RECON_MUSCL_VANLEER) return 2.0 * extrap * half_dq / (extrap + half_dq);
  return fabs(extrap) < fabs(half_dq) ? extrap : half_dq; /* minmod */
}

/* ComputeLeastSquaresGradients, src/operator_fluxes_ceed.c:993-1038, turned cell-centric: one thread per owned cell sums
 * weight * (q_right - q_left) over its interior edges in the reference's edge order (no atomics).  Ghost-cell gradients arrive
 * by halo exchange (CommunicateCellGradients, :1056-1098). */
__global__ void muscl_gradient_kernel(const SweArgs a, double *__restrict__ grad, int32_t n_owned) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_owned) return;
  const double q0 = a.h[i], q1 = a.hu[i], q2 = a.hv[i];
  double gx0 = 0.0, gx1 = 0.0, gx2 = 0.0, gy0 = 0.0, gy1 = 0.0, gy2 = 0.0;
#pragma unroll
  for (int k = 0; k < RDYB_ELL_WIDTH; ++k) {
    if (k >= a.width) break;
    const size_t q = (size_t)k * (size_t)a.stride + (size_t)i;
    const int nbr = a.nbr[q];
    if (nbr < 0) continue; /* empty or boundary slot */
    /* the weights carry the side (device_layout.c): left cell +w, right cell -w, both times (q_nbr - q_self).  Explicit FMAs in slot
     * order: K4M (swe_tiled_muscl_fixed.cuh) repeats exactly this sequence inside the tile kernel */
    const double d0 = a.h[nbr] - q0, d1 = a.hu[nbr] - q1, d2 = a.hv[nbr] - q2;
    const double cx = a.gcx[q], cy = a.gcy[q];
    gx0 = fma(cx, d0, gx0); gy0 = fma(cy, d0, gy0); gx1 = fma(cx, d1, gx1); gy1 = fma(cy, d1, gy1); gx2 = fma(cx, d2, gx2); gy2 = fma(cy, d2, gy2);
  }
  const size_t S = (size_t)a.stride;
  grad[i] = gx0; grad[S + i] = gx1; grad[2 * S + i] = gx2; grad[3 * S + i] = gy0; grad[4 * S + i] = gy1; grad[5 * S + i] = gy2;
}

__device__ __forceinline__ void courant_block_reduce(double cmax, long long carg, double *slot_val, long long *slot_arg) {
  /* max over the block; ties go to the smaller arg so the result does not depend on scheduling */
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) {
    const double ov = __shfl_down_sync(0xffffffffu, cmax, off);
    const long long oa = __shfl_down_sync(0xffffffffu, carg, off);
    if (ov > cmax || (ov == cmax && oa >= 0 && (carg < 0 || oa < carg))) { cmax = ov; carg = oa; }
  }
  __shared__ double sval[RDYB_BLOCK / 32];
  __shared__ long long sarg[RDYB_BLOCK / 32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) { sval[warp] = cmax; sarg[warp] = carg; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < RDYB_BLOCK / 32; ++w)
      if (sval[w] > cmax || (sval[w] == cmax && sarg[w] >= 0 && (carg < 0 || sarg[w] < carg))) { cmax = sval[w]; carg = sarg[w]; }
    if (cmax > *slot_val) { *slot_val = cmax; *slot_arg = carg; } /* the block owns its slot: plain read-modify-write */
  }
}

template <bool HR, int SRC, bool EULER, bool COURANT, int RECON = 0>
__global__ void __launch_bounds__(RDYB_BLOCK) swe_gather_kernel(const SweArgs a) {
  if (a.ctl && (!a.ctl->active || a.ctl->cur != a.ctl_buf)) return; /* device-side loop: not this launch's turn */
  const double dt_ = a.ctl ? a.ctl->dt_step : a.dt;
  const int i = a.begin + blockIdx.x * RDYB_BLOCK + threadIdx.x;
  double cmax = 0.0;
  long long carg = -1;
  if (i < a.end) {
    const double g = RDYB_GRAVITY;
    const double tiny = a.tiny_h;
    const double h = a.h[i], hu = a.hu[i], hv = a.hv[i];
    const double ia = a.inv_area[i];
    double u, v;
    swe_velocities<HR>(h, hu, hv, tiny, a.ha2, u, v);
    const double zc = HR ? a.zc[i] : 0.0;
    const double sq = swe_sqrt(h); /* HR re-evaluates it on the reconstructed depth */
    double F0 = 0.0, F1 = 0.0, F2 = 0.0;
#pragma unroll
    for (int k = 0; k < RDYB_ELL_WIDTH; ++k) {
      if (k >= a.width) break;
      const size_t q = (size_t)k * (size_t)a.stride + (size_t)i;
      const int nbr = a.nbr[q];
      if (nbr == RDYB_NO_EDGE) continue;
      const double sn = a.esn[q], cn = a.ecn[q], slen = a.eslen[q];
      const double scale = slen * ia; /* -L/A when this cell is the left cell, +L/A when it is the right cell */
      if (RECON != 0 && nbr >= 0) {
        /* second order: ReconstructFaceValues (src/operator_fluxes_ceed.c:1150-1195) + ApplyInteriorFlux2R (src/swe/swe_petsc.c:140-200) */
        const bool left = slen < 0.0;
        const size_t S = (size_t)a.stride;
        const int cl = left ? i : nbr, cr = left ? nbr : i;
        const double dxl = left ? a.dxs[q] : a.dxn[q], dyl = left ? a.dys[q] : a.dyn[q];
        const double dxr = left ? a.dxn[q] : a.dxs[q], dyr = left ? a.dyn[q] : a.dys[q];
        double fl[3], fr[3];
        const double *qv[3] = {a.h, a.hu, a.hv};
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          const double ql = qv[c][cl], qr = qv[c][cr];
          const double el = a.grad[c * S + cl] * dxl + a.grad[(3 + c) * S + cl] * dyl;
          const double er = a.grad[c * S + cr] * dxr + a.grad[(3 + c) * S + cr] * dyr;
          const double dq = qr - ql;
          fl[c] = ql + swe_limit_slope<RECON>(el, 0.5 * dq);
          fr[c] = qr + swe_limit_slope<RECON>(er, -0.5 * dq);
        }
        const double hl = fmax(0.0, fl[0]), hr = fmax(0.0, fr[0]);
        if (hl < tiny && hr < tiny) continue; /* src/swe/swe_petsc.c:179 */
        double ul, vl, ur, vr;
        swe_velocities<false>(hl, fl[1], fl[2], tiny, a.ha2, ul, vl);
        swe_velocities<false>(hr, fr[1], fr[2], tiny, a.ha2, ur, vr);
        const RoeFlux f = swe_roe_flux(hl, ul, vl, swe_sqrt(hl), hr, ur, vr, swe_sqrt(hr), sn, cn);
        if (COURANT) {
          const double cnum = f.amax * fabs(slen) * fmax(ia, a.inv_area[nbr]) * dt_;
          if (cnum > cmax) { cmax = cnum; carg = (long long)i * RDYB_ELL_WIDTH + k; }
        }
        F0 += f.f0 * scale;
        F1 += f.f1 * scale;
        F2 += f.f2 * scale;
      } else if (nbr >= 0) {
        const double hn = a.h[nbr];
        if (hn < tiny && h < tiny) continue; /* src/swe/swe_petsc.c:285, :1096 */
        const double hun = a.hu[nbr], hvn = a.hv[nbr];
        double un, vn;
        swe_velocities<HR>(hn, hun, hvn, tiny, a.ha2, un, vn);
        const bool left = slen < 0.0;
        double hl = left ? h : hn, ul = left ? u : un, vl = left ? v : vn;
        double hr = left ? hn : h, ur = left ? un : u, vr = left ? vn : v;
        bool do_flux = true;
        double c1 = 0.0, c2 = 0.0;
        if (HR) {
          /* hydrostatic reconstruction, src/swe/swe_petsc.c:1060-1075 */
          const double zn = a.zc[nbr];
          const double zl = left ? zc : zn, zr = left ? zn : zc;
          const double zmax = fmax(zl, zr);
          const double hls = fmax(0.0, (hl + zl) - zmax), hrs = fmax(0.0, (hr + zr) - zmax);
          /* pressure correction on this cell's side, outside the inner guard (:1137-1154) */
          const double hme = left ? hls : hrs;
          const double corr = 0.5 * g * (h * h - hme * hme);
          c1 = corr * cn * scale;
          c2 = corr * sn * scale;
          hl = hls;
          hr = hrs;
          do_flux = (hls > tiny || hrs > tiny);
        }
        if (do_flux) {
          double sql, sqr;
          if (HR) { sql = swe_sqrt(hl); sqr = swe_sqrt(hr); }
          else { const double sqn = swe_sqrt(hn); sql = left ? sq : sqn; sqr = left ? sqn : sq; }
          RoeFlux f = swe_roe_flux(hl, ul, vl, sql, hr, ur, vr, sqr, sn, cn);
          if (a.bs_eta) { const double bs = swe_bs2002_term(a, q); f.f1 += bs * cn; f.f2 += bs * sn; }
          if (COURANT) {
            const double cnum = f.amax * fabs(slen) * fmax(ia, a.inv_area[nbr]) * dt_; /* amax*L/min(A_l,A_r)*dt */
            if (cnum > cmax) { cmax = cnum; carg = (long long)i * RDYB_ELL_WIDTH + k; }
          }
          F0 += f.f0 * scale;
          F1 += f.f1 * scale;
          F2 += f.f2 * scale;
        }
        if (HR) { F1 += c1; F2 += c2; }
      } else {
        /* boundary edge: the cell is the left cell; ApplyBoundaryFlux uses the non-strict wet test for both operators */
        const int j = -(nbr + 1);
        double ul, vl;
        swe_velocities<false>(h, hu, hv, tiny, a.ha2, ul, vl);
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
        RoeFlux f = {0.0, 0.0, 0.0, 0.0};
        if (ok && !(h < tiny && hr < tiny)) {
          f = swe_roe_flux(h, ul, vl, swe_sqrt(h), hr, ur, vr, swe_sqrt(hr), sn, cn);
          if (a.bs_eta) { const double bs = swe_bs2002_term(a, q); f.f1 += bs * cn; f.f2 += bs * sn; }
          if (COURANT) {
            const double cnum = f.amax * fabs(slen) * ia * dt_;
            if (cnum > cmax) { cmax = cnum; carg = (long long)i * RDYB_ELL_WIDTH + k; }
          }
          F0 += f.f0 * scale;
          F1 += f.f1 * scale;
          F2 += f.f2 * scale;
        }
        if (a.bflux) { /* boundary_fluxes and boundary_fluxes_accum += dt*flux, src/swe/swe_petsc.c:623 */
          a.bflux[j] = f.f0; a.bflux[a.nbe + j] = f.f1; a.bflux[2 * a.nbe + j] = f.f2;
          a.bacc[j] += dt_ * f.f0; a.bacc[a.nbe + j] += dt_ * f.f1; a.bacc[2 * a.nbe + j] += dt_ * f.f2;
        }
      }
    }
    if (a.fdiv) { a.fdiv[i] = F0; a.fdiv[a.stride + i] = F1; a.fdiv[2 * a.stride + i] = F2; }
    /* sources */
    double bedx = 0.0, bedy = 0.0, tbx = 0.0, tby = 0.0;
    if (!HR) { bedx = a.dzdx[i] * g * h; bedy = a.dzdy[i] * g * h; }
    if (h >= tiny) swe_friction<SRC>(h, hu, hv, F1, F2, bedx, bedy, a.mann[i], dt_, a.xq_thr, tbx, tby);
    const double S1 = a.s1 ? a.s1[i] : 0.0, S2 = a.s2 ? a.s2[i] : 0.0;
    F0 += a.s0[i];
    F1 += -bedx - tbx + S1;
    F2 += -bedy - tby + S2;
    if (a.prim) {
      const bool wet = h >= tiny;
      const double r = h * swe_rcp(h * h + a.ha2);
      a.prim[i] = h; a.prim[a.stride + i] = wet ? hu * r : 0.0; a.prim[2 * a.stride + i] = wet ? hv * r : 0.0;
    }
    if (EULER) {
      a.o0[i] = h + dt_ * F0;
      a.o1[i] = hu + dt_ * F1;
      a.o2[i] = hv + dt_ * F2;
    } else {
      a.o0[i] = F0;
      a.o1[i] = F1;
      a.o2[i] = F2;
    }
  }
  if (COURANT && a.cmax) courant_block_reduce(cmax, carg, &a.cmax[blockIdx.x], &a.carg[blockIdx.x]);
}

/* ---- layout conversion between the reference's AoS Vecs and the device SoA ---- */
/* dst SoA[c*stride + d] = src AoS[3*map[d] + c] for d in [0,n) */
__global__ void aos_to_soa_kernel(const double *__restrict__ src, const int32_t *__restrict__ map, double *__restrict__ dst, int32_t n, int32_t stride) {
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d < n) {
    const size_t s = 3 * (size_t)map[d];
    dst[d] = src[s]; dst[stride + d] = src[s + 1]; dst[2 * stride + d] = src[s + 2];
  }
}
/* dst AoS[3*map[d] + c] = src SoA[c*stride + d] */
__global__ void soa_to_aos_kernel(const double *__restrict__ src, const int32_t *__restrict__ map, double *__restrict__ dst, int32_t n, int32_t stride) {
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d < n) {
    const size_t s = 3 * (size_t)map[d];
    dst[s] = src[d]; dst[s + 1] = src[stride + d]; dst[s + 2] = src[2 * stride + d];
  }
}
/* the same conversions walked in the HOST order (chunks of the pipelined host-Vec apply):
 * dst SoA[c*stride + map[i]] = src AoS[3*i + c] (map[i] < 0: a view cell the layout dropped) and dst AoS[3*i + c] = src SoA[c*stride + map[i]] */
__global__ void aos_scatter_soa_kernel(const double *__restrict__ src, const int32_t *__restrict__ map, double *__restrict__ dst, int32_t n, int32_t stride) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const int d = map[i];
    if (d >= 0) { dst[d] = src[3 * (size_t)i]; dst[stride + d] = src[3 * (size_t)i + 1]; dst[2 * stride + d] = src[3 * (size_t)i + 2]; }
  }
}
__global__ void soa_gather_aos_kernel(const double *__restrict__ src, const int32_t *__restrict__ map, double *__restrict__ dst, int32_t n, int32_t stride) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const int d = map[i];
    dst[3 * (size_t)i] = src[d]; dst[3 * (size_t)i + 1] = src[stride + d]; dst[3 * (size_t)i + 2] = src[2 * stride + d];
  }
}
/* dst[d] = src[map[d]] / dst[map[d]] = src[d] */
__global__ void gather_kernel(const double *__restrict__ src, const int32_t *__restrict__ map, double *__restrict__ dst, int32_t n) {
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d < n) dst[d] = src[map[d]];
}
__global__ void scatter_kernel(const double *__restrict__ src, const int32_t *__restrict__ map, double *__restrict__ dst, int32_t n) {
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d < n) dst[map[d]] = src[d];
}

/* halo pack: buf[c*nsend + k] = state[c*stride + cells[k]] (replaces the PetscSF pack of DMGlobalToLocal, src/rdysetup.c:1133) */
__global__ void halo_pack_kernel(const double *__restrict__ state, const int32_t *__restrict__ cells, double *__restrict__ buf, int32_t nsend,
                                 int32_t stride) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < nsend) {
    const int c = cells[k];
    buf[k] = state[c]; buf[nsend + k] = state[stride + c]; buf[2 * nsend + k] = state[2 * stride + c];
  }
}
/* halo unpack: ghosts of all peers are the contiguous device cells [n_owned, n_owned+nrecv) */
__global__ void halo_unpack_kernel(const double *__restrict__ buf, double *__restrict__ state, int32_t nrecv, int32_t n_owned, int32_t stride) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k < nrecv) {
    state[n_owned + k] = buf[k]; state[stride + n_owned + k] = buf[nrecv + k]; state[2 * stride + n_owned + k] = buf[2 * nrecv + k];
  }
}

/* ---- peer-memory halo (NVLink P2P): the owner writes its border cells straight into the neighbours' ghost slots ----
 * Replaces DMGlobalToLocalBegin/End (src/rdysetup.c:1133-1134) on the Euler fast path: no pack buffer, no NCCL kernel, no
 * unpack; one small kernel per step stores 24 B per shared-face cell over NVLink and then raises a step counter in each
 * neighbour's flag word; the neighbour's stream waits for that counter before its frontier cells run. */
#define RDYB_MAX_PEERS 8
struct PeerTable {
  double *state[RDYB_MAX_PEERS][3];   /* neighbour's two state buffers (ping-pong) and its RK stage buffer, mapped through CUDA IPC */
  unsigned int *flag[RDYB_MAX_PEERS]; /* the word in the neighbour's flag array that belongs to this rank */
  int32_t stride[RDYB_MAX_PEERS];     /* neighbour's leading dimension */
  int32_t ghost0[RDYB_MAX_PEERS];     /* neighbour's device index of the first ghost cell it receives from this rank */
  int32_t send_begin[RDYB_MAX_PEERS + 1];
  int32_t num_peers;
};
__global__ void __launch_bounds__(1024) halo_push_kernel(const double *__restrict__ state, int32_t stride, const int32_t *__restrict__ send_cells,
                                                         const PeerTable tbl, int32_t buf, unsigned int epoch) {
  for (int p = 0; p < tbl.num_peers; ++p) {
    double *dst = tbl.state[p][buf];
    const int32_t rs = tbl.stride[p], g0 = tbl.ghost0[p], b = tbl.send_begin[p], n = tbl.send_begin[p + 1] - b;
    for (int k = threadIdx.x; k < n; k += blockDim.x) {
      const int c = send_cells[b + k];
      dst[g0 + k] = state[c]; dst[(size_t)rs + g0 + k] = state[(size_t)stride + c]; dst[2 * (size_t)rs + g0 + k] = state[2 * (size_t)stride + c];
    }
  }
  __threadfence_system(); /* the stores above are visible system-wide before the counter */
  __syncthreads();
  if (threadIdx.x < tbl.num_peers) {
    *reinterpret_cast<volatile unsigned int *>(tbl.flag[threadIdx.x]) = epoch;
    __threadfence_system();
  }
}
/* wait until every neighbour has raised its counter in my flag array to `epoch`; gives up after ~timeout_cycles and
 * records the failure in *status (a hung peer must not hang this GPU) */
__global__ void halo_wait_kernel(const unsigned int *flags, int32_t num_peers, unsigned int epoch, long long timeout_cycles, int *status) {
  if (threadIdx.x < num_peers) {
    const volatile unsigned int *f = flags + threadIdx.x;
    const long long t0 = clock64();
    while ((int)(*f - epoch) < 0) {
      __nanosleep(100);
      if (clock64() - t0 > timeout_cycles) { atomicExch(status, 1 + threadIdx.x); break; }
    }
  }
  __threadfence_system();
}

/* y[c*stride+i] = x[c*stride+i] + w*k[c*stride+i], i < n (TSRK4 stage vectors) */
__global__ void stage_axpy_kernel(const double *__restrict__ x, const double *__restrict__ k, double w, double *__restrict__ y, int32_t n, int32_t stride) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
#pragma unroll
    for (int c = 0; c < 3; ++c) y[c * (size_t)stride + i] = x[c * (size_t)stride + i] + w * k[c * (size_t)stride + i];
  }
}
/* u += w0*k0, then += w1*k1, ... in that order (VecMAXPY-like, sequential per entry) */
__global__ void rk4_combine_kernel(double *__restrict__ u, const double *__restrict__ k0, const double *__restrict__ k1, const double *__restrict__ k2,
                                   const double *__restrict__ k3, double w0, double w1, double w2, double w3, int32_t n, int32_t stride) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const size_t q = c * (size_t)stride + i;
      double x = u[q];
      x += w0 * k0[q]; x += w1 * k1[q]; x += w2 * k2[q]; x += w3 * k3[q];
      u[q] = x;
    }
  }
}

/* deterministic two-stage sum of area*h over owned cells: fixed blocks, fixed tree */
__global__ void mass_partial_kernel(const double *__restrict__ h, const double *__restrict__ area, int32_t n, double *__restrict__ partial) {
  __shared__ double s[256];
  double acc = 0.0;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) acc += area[i] * h[i];
  s[threadIdx.x] = acc;
  __syncthreads();
  for (int off = 128; off > 0; off >>= 1) {
    if ((int)threadIdx.x < off) s[threadIdx.x] += s[threadIdx.x + off];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = s[0];
}

#endif
