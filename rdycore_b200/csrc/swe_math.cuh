/* swe_math.cuh — fp64 device arithmetic of the SWE right-hand side (sm_100a).
 *
 * Each function states the reference code it reproduces (paths relative to the RDycore tree).  The arithmetic is
 * the reference's, re-associated only where noted: sqrt() replaces pow(x,0.5), rcbrt() replaces pow(h,-1/3),
 * reciprocals are shared, and the compiler may contract a*b+c into FMA.  Every such change moves a result by a few
 * ulp; the parity tests bound the accumulated effect (<= 1e-11 relative after 1000 steps).
 */
#ifndef RDYB_SWE_MATH_CUH
#define RDYB_SWE_MATH_CUH

#define RDYB_GRAVITY 9.806 /* src/swe/swe_types_petsc.h:7 */

#define RDYB_SQRT_GRAVITY 3.1314533367112465 /* sqrt of the double 9.806 */

/* ---- branch-free fp64 reciprocal / rsqrt / sqrt ----
 * MUFU seed (rcp.approx / rsqrt.approx, ~2^-20) + two Newton steps -> <= ~1 ulp, without the denormal / inf / NaN slow
 * path that the IEEE-rounded `1.0/x` and `sqrt(x)` sequences carry (a CALL plus ~10 integer instructions each).  Valid
 * for normal, positive, finite arguments, which is all this path produces (depths in [tiny_h, ~1e4], areas, speeds). */
__device__ __forceinline__ double swe_rcp(double a) {
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
  double e = fma(-a, y, 1.0);
  y = fma(y, e, y);
  e = fma(-a, y, 1.0);
  y = fma(y, e, y);
  return y;
}
/* rsqrt.approx.ftz flushes a subnormal argument to +0 and returns +Inf, and a Newton step on Inf gives NaN.  Squared speeds
 * u*u + v*v DO pass through the subnormal range: the leading edge of a disturbance entering still water decays geometrically
 * and crosses |u| ~ 1e-162..1e-154 (the reference's sqrt() returns a tiny finite value there).  swe_sqrt therefore returns 0
 * for every argument below 2^-1000 (true value < 1e-150: an absolute error far below one ulp of anything it is added to or
 * multiplied into on this path) - the same select the a == 0 case always needed, so the hot path pays nothing. */
#define RDYB_TINY_ARG 9.3326361850321888e-302 /* 2^-1000 */
/* 1/sqrt(a) for normal positive a */
__device__ __forceinline__ double swe_rsqrt_core(double a) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
  const double ha = 0.5 * a;
  double e = fma(-ha, y * y, 0.5);
  y = fma(y, e, y);
  e = fma(-ha, y * y, 0.5);
  y = fma(y, e, y);
  return y;
}
/* sqrt(a) for a >= 0; exactly 0 for a < 2^-1000 (0 and the subnormal range included); one Heron correction after a*rsqrt(a) */
__device__ __forceinline__ double swe_sqrt(double a) {
  const double y = swe_rsqrt_core(a);
  double s = a * y;
  const double r = fma(-s, s, a);
  s = fma(r, 0.5 * y, s);
  return a >= RDYB_TINY_ARG ? s : 0.0;
}

/* sqrt(a) = a * rsqrt(a) without the Heron correction: <= ~1.5 ulp instead of <= ~1 ulp, three fp64 instructions fewer.  Used per
 * edge side in K4's edge pass, where sqrt(h) only enters the Roe averages and the celerities (an error of one ulp there is the
 * effect of a 2-ulp change of h). */
__device__ __forceinline__ double swe_sqrt_fast(double a) {
  const double y = swe_rsqrt_core(a >= RDYB_TINY_ARG ? a : 1.0);
  return a >= RDYB_TINY_ARG ? a * y : 0.0;
}

/* a^(-1/3) for normal positive a (depths in [tiny_h, ~1e4]): fp32 MUFU seed exp2(-log2(a)/3) (~2^-21) + two Newton steps
 * y <- y + y (1 - a y^3) / 3 (error e -> 2 e^2), ~15 instructions instead of the ~40 of the library rcbrt() with its
 * exponent-splitting and special-case paths */
__device__ __forceinline__ double swe_rcbrt(double a) {
  double y = (double)exp2f(-0.333333343f * __log2f((float)a));
#pragma unroll
  for (int it = 0; it < 2; ++it) {
    const double e = fma(-(a * y), y * y, 1.0);
    y = fma(y * e, 0.33333333333333331, y);
  }
  return y;
}

struct RoeFlux {
  double f0, f1, f2, amax;
};

/* ComputeRiemannVelocities, src/swe/swe_petsc.c:57-73: u = hu*h/(h^2+h_a^2), 0 when h < tiny_h.
 * STRICT selects the HR operator's variant, wet iff h > tiny_h (src/swe/swe_petsc.c:1077-1092). */
template <bool STRICT>
__device__ __forceinline__ void swe_velocities(double h, double hu, double hv, double tiny_h, double ha2, double &u, double &v) {
  const bool wet = STRICT ? (h > tiny_h) : !(h < tiny_h);
  if (wet) {
    /* h_a = 0 (the default): hu h / h^2 = hu / h, one reciprocal of h instead of the product chain (a uniform branch) */
    const double r = ha2 == 0.0 ? swe_rcp(h) : h * swe_rcp(h * h + ha2);
    u = hu * r;
    v = hv * r;
  } else {
    u = 0.0;
    v = 0.0;
  }
}

/* Primitive state of one edge side, computed where the edge needs it (K4's fused variant: no per-cell pass, no u/v/sqrt(h)
 * scratch, one CTA barrier per tile less): u, v as swe_velocities<false> and sq = sqrt(h), from ONE MUFU seed.  With
 * y = 1/sqrt(h) (two Newton steps, ~1 ulp): sqrt(h) = h y and 1/h = y y, ~1.5 and ~2.5 ulp - no correction steps: the parity
 * tests bound the effect (it is far below the difference between pow(x, .5) on the CPU and any sqrt here).  The general
 * regularised form hu h / (h^2 + h_a^2) is kept for h_a != 0 (a uniform branch). */
__device__ __forceinline__ void swe_edge_side(double h, double hu, double hv, double tiny_h, double ha2, double &u, double &v, double &sq) {
  const bool pos = h >= RDYB_TINY_ARG;
  const double hs = pos ? h : 1.0;
  const double y = swe_rsqrt_core(hs);
  sq = pos ? hs * y : 0.0;
  double inv;
  if (ha2 == 0.0) inv = y * y; /* 1/h to ~2.5 ulp */
  else inv = hs * swe_rcp(hs * hs + ha2);
  const bool wet = !(h < tiny_h);
  u = wet ? hu * inv : 0.0;
  v = wet ? hv * inv : 0.0;
}

/* ComputeSWERoeEigenspectrum + the flux loop of ComputeSWERoeFlux, src/swe/swe_roe_flux_petsc.h:15-81, 108-128.
 * Inputs are (h,u,v,sqrt h) of the left and right state and the edge normal (cn,sn).  At least one of hl,hr must be > 0.
 * Re-association vs the reference: sqrt(g h) = sqrt(g) sqrt(h); chat and 1/chat from one rsqrt; one shared 1/(duml+dumr); the
 * pressure terms and the w1/w3 columns of R |Lambda| dW are summed before they meet the normal (fewer multiplications). */
template <bool FIX>
__device__ __forceinline__ RoeFlux swe_roe_flux_t(double hl, double ul, double vl, double sql, double hr, double ur, double vr, double sqr, double sn,
                                                  double cn, bool &needfix) {
  /* sql = sqrt(hl), sqr = sqrt(hr) come from the caller (once per cell, not once per edge) */
  const double g = RDYB_GRAVITY;
  const double duml = sql, dumr = sqr;
  const double cl = RDYB_SQRT_GRAVITY * sql, cr = RDYB_SQRT_GRAVITY * sqr; /* sqrt(g*h) */
  const double hhat = duml * dumr;
  const double rsum = swe_rcp(duml + dumr);
  /* (duml ul + dumr ur) / (duml + dumr) = ul + w (ur - ul) with w = dumr / (duml + dumr): one FMA per component */
  const double wr = dumr * rsum;
  const double uhat = fma(wr, ur - ul, ul);
  const double vhat = fma(wr, vr - vl, vl);
  const double c2 = 0.5 * g * (hl + hr);
  const double rchat = swe_rsqrt_core(c2); /* c2 >= g tiny_h / 2: normal */
  const double chat = c2 * rchat;
  const double uperp = uhat * cn + vhat * sn;
  const double dh = hr - hl, du = ur - ul, dv = vr - vl;
  const double dupar = -du * sn + dv * cn;
  const double duperp = du * cn + dv * sn;
  const double uperpl = ul * cn + vl * sn, uperpr = ur * cn + vr * sn;
  double a1 = fabs(uperp - chat);
  const double a2 = fabs(uperp);
  double a3 = fabs(uperp + chat);
  /* critical flow fix, :55-69: 2 ((uperpr -+ cr) - (uperpl -+ cl)) = 2 (d(uperp) -+ d(c)), with d(uperp) = duperp */
  const double dc = cr - cl;
  const double da1 = fmax(0.0, 2.0 * (duperp - dc));
  const double da3 = fmax(0.0, 2.0 * (duperp + dc));
  if (FIX) {
    /* a subnormal da (rcp.approx.ftz would return Inf, and 0 * Inf = NaN) is clamped to 2^-1000: a < da makes a * a underflow to
     * exactly 0 there, so the quotient is the 0 the reference's division gives */
    if (a1 < da1) a1 = 0.5 * (a1 * a1 * swe_rcp(fmax(da1, RDYB_TINY_ARG)) + da1);
    if (a3 < da3) a3 = 0.5 * (a3 * a3 * swe_rcp(fmax(da3, RDYB_TINY_ARG)) + da3);
    needfix = false;
  } else {
    needfix = (a1 < da1) || (a3 < da3); /* the caller re-evaluates these (rare) edges with FIX = true */
  }
  const double t = hhat * duperp * rchat;
  const double w1 = a1 * (0.5 * (dh - t)), w2 = a2 * (hhat * dupar), w3 = a3 * (0.5 * (dh + t));
  const double psum = 0.5 * g * (hl * hl + hr * hr); /* pl + pr */
  const double ql = uperpl * hl, qr = uperpr * hr;
  RoeFlux F;
  /* F = 0.5*(FL + FR - R |Lambda| dW) with R = [[1,0,1],[uhat-chat*cn,-sn,uhat+chat*cn],[vhat-chat*sn,cn,vhat+chat*sn]];
   * the w1 and w3 columns are summed as uhat (w1 + w3) + chat cn (w3 - w1) (and the same with vhat, sn) */
  const double ws = w1 + w3, wd = chat * (w3 - w1);
  F.f0 = 0.5 * (ql + qr - ws);
  F.f1 = 0.5 * (ul * ql + ur * qr + psum * cn - uhat * ws - wd * cn + sn * w2);
  F.f2 = 0.5 * (vl * ql + vr * qr + psum * sn - vhat * ws - wd * sn - cn * w2);
  F.amax = chat + fabs(uperp);
  return F;
}
__device__ __forceinline__ RoeFlux swe_roe_flux(double hl, double ul, double vl, double sql, double hr, double ur, double vr, double sqr, double sn,
                                                double cn) {
  bool unused;
  return swe_roe_flux_t<true>(hl, ul, vl, sql, hr, ur, vr, sqr, sn, cn, unused);
}

/* ApplyReflectingBC, src/swe/swe_petsc.c:434-461 (acts on velocities) */
__device__ __forceinline__ void swe_reflecting_bc(double hl, double ul, double vl, double sn, double cn, double &hr, double &ur, double &vr) {
  const double dum1 = sn * sn - cn * cn, dum2 = 2.0 * sn * cn;
  hr = hl;
  ur = ul * dum1 - vl * dum2;
  vr = -ul * dum2 - vl * dum1;
}

/* ApplyCriticalOutflowBC, src/swe/swe_petsc.c:465-503.  Returns false when the edge carries no flux (inflow: both
 * sides are set dry by the reference, and the dry-dry guard at :593 then skips it). */
__device__ __forceinline__ bool swe_critical_outflow_bc(double hl, double ul, double vl, double sn, double cn, double &hr, double &ur, double &vr) {
  const double uperp = ul * cn + vl * sn;
  if (uperp < 0.0) return false;
  const double q = hl * fabs(uperp);
  hr = cbrt(q * q / RDYB_GRAVITY);
  const double vel = swe_sqrt(RDYB_GRAVITY * hr);
  ur = vel * cn;
  vr = vel * sn;
  return true;
}

/* Bed friction.  METHOD 0: ApplySourceSemiImplicit src/swe/swe_petsc.c:758-785; METHOD 1: ApplySourceImplicitXQ2018
 * src/swe/swe_petsc.c:870-912.  (Fx,Fy) is the flux divergence of this cell, (bedx,bedy) the bed-slope term. */
template <int METHOD>
__device__ __forceinline__ void swe_friction(double h, double hu, double hv, double Fx, double Fy, double bedx, double bedy, double n, double dt,
                                             double xq_thr, double &tbx, double &tby) {
  const double g = RDYB_GRAVITY;
  const double rh = swe_rcp(h);
  const double rc = swe_rcbrt(h); /* h^(-1/3) */
  if (METHOD == 0) {
    const double u = hu * rh, v = hv * rh;
    const double Cd = g * (n * n) * rc;
    const double vel = swe_sqrt(u * u + v * v);
    const double tb = Cd * vel * rh;
    const double factor = tb * swe_rcp(1.0 + dt * tb);
    tbx = (hu + dt * Fx - dt * bedx) * factor;
    tby = (hv + dt * Fy - dt * bedy) * factor;
  } else {
    const double Ax = Fx - bedx, Ay = Fy - bedy;
    const double mx = hu + Ax * dt, my = hv + Ay * dt;
    const double rc2 = rc * rc, rc4 = rc2 * rc2; /* h^(-4/3) */
    const double mxh = mx * rh, myh = my * rh;
    const double gn2 = g * (n * n);
    const double lambda = gn2 * rc4 * swe_sqrt(mxh * mxh + myh * myh);
    double qx, qy;
    if (dt * lambda < xq_thr) {
      qx = mx;
      qy = my;
    } else {
      const double s = swe_sqrt(1.0 + 4.0 * dt * lambda);
      const double rd = -swe_rcp(2.0 * dt * lambda);
      qx = (mx - mx * s) * rd;
      qy = (my - my * s) * rd;
    }
    const double qmag = swe_sqrt(qx * qx + qy * qy);
    const double c = gn2 * (rc4 * rc2 * rc) * qmag; /* g n^2 h^(-7/3) |q| */
    tbx = c * qx;
    tby = c * qy;
  }
}


/* The same friction terms in two halves, for kernels that overlap a barrier wait with the part that does not need the flux
 * divergence (K4): swe_friction_pre runs before the edge fluxes of the tile are complete, swe_friction_post after.  Operation
 * for operation the arithmetic of swe_friction, so the two forms give the same bits.
 * METHOD 0: p0 = tb / (1 + dt tb) (the whole nonlinear part).  METHOD 1: p0 = 1/h, p1 = h^(-1/3) (lambda needs the fluxes). */
template <int METHOD>
__device__ __forceinline__ void swe_friction_pre(double h, double hu, double hv, double n, double dt, double &p0, double &p1) {
  const double g = RDYB_GRAVITY;
  const double rh = swe_rcp(h);
  const double rc = swe_rcbrt(h); /* h^(-1/3) */
  if (METHOD == 0) {
    const double u = hu * rh, v = hv * rh;
    const double Cd = g * (n * n) * rc;
    const double vel = swe_sqrt(u * u + v * v);
    const double tb = Cd * vel * rh;
    p0 = tb * swe_rcp(1.0 + dt * tb);
    p1 = 0.0;
  } else {
    p0 = rh;
    p1 = rc;
  }
}
template <int METHOD>
__device__ __forceinline__ void swe_friction_post(double p0, double p1, double hu, double hv, double Fx, double Fy, double bedx, double bedy, double n,
                                                  double dt, double xq_thr, double &tbx, double &tby) {
  const double g = RDYB_GRAVITY;
  if (METHOD == 0) {
    tbx = (hu + dt * Fx - dt * bedx) * p0;
    tby = (hv + dt * Fy - dt * bedy) * p0;
  } else {
    const double rh = p0, rc = p1;
    const double Ax = Fx - bedx, Ay = Fy - bedy;
    const double mx = hu + Ax * dt, my = hv + Ay * dt;
    const double rc2 = rc * rc, rc4 = rc2 * rc2; /* h^(-4/3) */
    const double mxh = mx * rh, myh = my * rh;
    const double gn2 = g * (n * n);
    const double lambda = gn2 * rc4 * swe_sqrt(mxh * mxh + myh * myh);
    double qx, qy;
    if (dt * lambda < xq_thr) {
      qx = mx;
      qy = my;
    } else {
      const double s = swe_sqrt(1.0 + 4.0 * dt * lambda);
      const double rd = -swe_rcp(2.0 * dt * lambda);
      qx = (mx - mx * s) * rd;
      qy = (my - my * s) * rd;
    }
    const double qmag = swe_sqrt(qx * qx + qy * qy);
    const double c = gn2 * (rc4 * rc2 * rc) * qmag; /* g n^2 h^(-7/3) |q| */
    tbx = c * qx;
    tby = c * qy;
  }
}

#endif
