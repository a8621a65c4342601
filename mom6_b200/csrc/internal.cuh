// Internal definitions shared by the kernels and the plan (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cfloat>

// Padded HBM layout shared by every field of the plan (h, u and v points):
//   column of Fortran index i  : ci = i + coff   (cell i and its east face I=i share ci)
//   row    of Fortran index j  : rj = j + roff   (cell j and its north face J=j share rj)
//   element (i,j,k)            : ci + P*rj + plane*(k-1)
// coff is chosen so that the first computational cell (isc) sits on a multiple
// of 16 doubles (128 B) and P is a multiple of 16, so every row starts 128-byte
// aligned.  Column/row for index isd-1 / jsd-1 always exist (symmetric memory).
struct Geom {
  int P, R;            // pitch (columns) and rows of one padded plane
  long long plane;     // P*R
  int nk;
  int coff, roff;
  int isd, ied, jsd, jed;      // data domain (Fortran indices)
  int isc, iec, jsc, jec;      // computational domain
  int IsdB, JsdB;
};

struct Range { int is, ie, js, je; };   // Fortran-index cell range of a sweep

// Per-tracer device record (SoA field pointers)
struct TracerDev {
  double *buf[2];        // ping-pong storage; cur selects the live one
  double *ad_x, *ad_y, *ad2d_x, *ad2d_y, *adv_xy;
  double conc_underflow;
};

// Tracer table passed by value to the streaming kernels (constant bank).
#define MOM6ADV_MAXT 64
struct TrTable {
  const double *src[MOM6ADV_MAXT];
  double *dst[MOM6ADV_MAXT];
  double cu[MOM6ADV_MAXT];       // conc_underflow
  int n;
};

// Flux-diagnostic pointers of the tracers of a TrTable (same order); nullptr = unassociated.
//   ad: ad_x or ad_y of the sweep's direction (:665-667 / :1070-1076), axy: advection_xy (:671-676 / :1051-1056)
struct DiagTable {
  double *ad[MOM6ADV_MAXT];
  double *axy[MOM6ADV_MAXT];
};

// tracer_hordiff's per-call scalars (MOM_tracer_hor_diff.F90:380-386, :543-550)
struct HdParams {
  double I_numitts, ml_scale, h_neglect;
  int diffuse_ml, nkml, nk_rho_varies;
};

// diffusive-flux diagnostics of tracer_hordiff for the tracers of a TrTable (same order); nullptr = unassociated
struct DfTable {
  double *dfx[MOM6ADV_MAXT], *dfy[MOM6ADV_MAXT];      // df_x, df_y     (MOM_tracer_hor_diff.F90:576-583)
  double *d2x[MOM6ADV_MAXT], *d2y[MOM6ADV_MAXT];      // df2d_x, df2d_y (:584-591)
};

// ---- gfortran intrinsic semantics (argument-order compare-select) -----------
__device__ __forceinline__ double gmax2(double a, double b) { return (b > a) ? b : a; }
__device__ __forceinline__ double gmin2(double a, double b) { return (b < a) ? b : a; }
__device__ __forceinline__ double gmax3(double a, double b, double c) { double m = (b > a) ? b : a; return (c > m) ? c : m; }
__device__ __forceinline__ double gmin3(double a, double b, double c) { double m = (b < a) ? b : a; return (c < m) ? c : m; }

// x/3 and x/6, correctly rounded (bit-identical to IEEE division) without the
// ~30-instruction division sequence: q0 = RN(x*RN(1/d)), exact residual by FMA,
// one correction step (Markstein).  Falls back to true division outside the
// range where the residual is guaranteed exact.
__device__ __forceinline__ double div3(double x) {
  double ax = fabs(x);
  if (ax > 1e-280 && ax < 1e300) {
    const double c = 0.33333333333333331;   // RN(1/3)
    double q = __dmul_rn(x, c);
    double r = __fma_rn(-3.0, q, x);
    return __fma_rn(r, c, q);
  }
  return x / 3.0;
}
__device__ __forceinline__ double div6(double x) {
  double ax = fabs(x);
  if (ax > 1e-280 && ax < 1e300) {
    const double c = 0.16666666666666666;   // RN(1/6)
    double q = __dmul_rn(x, c);
    double r = __fma_rn(-6.0, q, x);
    return __fma_rn(r, c, q);
  }
  return x / 6.0;
}

// PLM slope, MOM_tracer_advect.F90:427-431 / :809-813; mm = mask(I)*mask(I-1)
__device__ __forceinline__ double plm_slope(double Tp, double Tc, double Tm, double mm) {
  double dMx = gmax3(Tp, Tc, Tm) - Tc;
  double dMn = Tc - gmin3(Tp, Tc, Tm);
  double d = Tp - Tm;
  double s = gmin3(0.5 * fabs(d), 2.0 * dMx, 2.0 * dMn);
  return mm * copysign(s, d);
}

enum { SCHEME_PLM = 0, SCHEME_PPM_CW = 1, SCHEME_PPM_H3 = 2 };

// PPM edge values + limiter for the upstream cell, :526-548 / :911-933.
//   t = (T[c-2], T[c-1], T[c], T[c+1], T[c+2]);  mm = (mm[c-1], mm[c], mm[c+1])
template <int SCHEME>
__device__ __forceinline__ void ppm_parabola(double Tm2, double Tm, double Tc, double Tp, double Tp2,
                                             double mmm, double mmc, double mmp,
                                             double &aL, double &aR, double &a6) {
  if (SCHEME == SCHEME_PPM_H3) {
    aL = div6(5. * Tc + (2. * Tm - Tp));
    aL = gmax2(gmin2(Tc, Tm), aL); aL = gmin2(gmax2(Tc, Tm), aL);
    aR = div6(5. * Tc + (2. * Tp - Tm));
    aR = gmax2(gmin2(Tc, Tp), aR); aR = gmin2(gmax2(Tc, Tp), aR);
  } else {
    double sm = plm_slope(Tc, Tm, Tm2, mmm);
    double sc = plm_slope(Tp, Tc, Tm, mmc);
    double sp = plm_slope(Tp2, Tp, Tc, mmp);
    aL = 0.5 * ((Tm + Tc) + div3(sm - sc));
    aR = 0.5 * ((Tc + Tp) + div3(sc - sp));
  }
  double dA = aR - aL, mA = 0.5 * (aR + aL);
  double t1 = dA * (Tc - mA), t2 = div6(dA * dA);
  if (mmc * (Tp - Tc) * (Tc - Tm) <= 0.) { aL = Tc; aR = Tc; }
  else if (t1 > t2) { aL = 3. * Tc - 2. * aR; }
  else if (t1 < -t2) { aR = 3. * Tc - 2. * aL; }
  a6 = 6. * Tc - 3. * (aR + aL);
}

// Tracer flux through the face between cells a and b=a+1 (:517-577 / :902-963).
//   t[0..5] = T[a-2..b+2],  mm[0..3] = mm[a-1], mm[a], mm[b], mm[b+1]
template <int SCHEME>
__device__ __forceinline__ double face_flux(const double t[6], const double mm[4], double uhh, double cfl) {
  if (SCHEME == SCHEME_PLM) {
    if (uhh >= 0.0) {
      double s = plm_slope(t[3], t[2], t[1], mm[1]);
      return uhh * (t[2] + 0.5 * s * (1. - cfl));
    } else {
      double s = plm_slope(t[4], t[3], t[2], mm[2]);
      return uhh * (t[3] - 0.5 * s * (1. - cfl));
    }
  } else {
    double aL, aR, a6;
    const double c23 = 2. / 3.;
    if (uhh >= 0.0) {
      ppm_parabola<SCHEME>(t[0], t[1], t[2], t[3], t[4], mm[0], mm[1], mm[2], aL, aR, a6);
      return uhh * (aR - 0.5 * cfl * ((aR - aL) - a6 * (1. - c23 * cfl)));
    } else {
      ppm_parabola<SCHEME>(t[1], t[2], t[3], t[4], t[5], mm[1], mm[2], mm[3], aL, aR, a6);
      return uhh * (aL + 0.5 * cfl * ((aR - aL) + a6 * (1. - c23 * cfl)));
    }
  }
}
