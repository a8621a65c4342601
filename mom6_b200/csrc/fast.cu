// Streaming tracer-sweep kernels (the production path when no flux diagnostic
// is associated).  Same arithmetic as kernels.cu / the reference
// (/root/reference/src/tracer/MOM_tracer_advect.F90:517-577,652-662 for x,
// :902-963,1043-1047 for y), restructured so that every slope, edge value,
// parabola and face flux is computed exactly ONCE per cell / face:
//
//   k_x_fast  one warp = 128 consecutive cells of one row (4 per lane, LDG.128),
//             all tracers looped in-thread with the tracer-independent mass
//             quantities (uhh, CFL, hlst, Ihnew, mask product) held in
//             registers; neighbour slopes / edges / outflows cross lanes by
//             warp shuffle.  Out of place (buf[src] -> buf[1-src]); tiles
//             overlap by 8 columns (120 useful of 128).
//   k_y_fast  one thread = two adjacent columns of one tracer, marching in j
//             with the whole stencil pipeline in registers (no shared memory,
//             no recomputation, in place); the warps of a CTA are different
//             tracers of the same columns, so the mass rows are L1 hits for
//             all but the first warp.
//
// Exact-by-construction rewrites used:
//   * max(Tp,Tc,Tm)-Tc and Tc-min(Tp,Tc,Tm) are obtained from the face
//     differences a_hi = Tp-Tc and a_lo = Tc-Tm (rounding is monotone and
//     sign-symmetric), so the PLM slope magnitude is
//     min(0.5|Tp-Tm|, 2|a_hi|, 2|a_lo|) when a_hi and a_lo have the same sign
//     and 0 otherwise; its sign is overwritten by sign(.., Tp-Tm) in the
//     reference, so zero signs cannot differ.
//   * min over non-negative doubles is an integer compare of the bit patterns.
//   * the CW84 edge value aR(i) and aL(i+1) are the same expression.
//   * a face's flux is evaluated by its upstream cell only.
#include "internal.cuh"
#include "kernels.h"

// min of two non-negative (sign bit clear) doubles: integer compare of the bit patterns
__device__ __forceinline__ double minpos(double x, double y) {
  return (__double_as_longlong(y) < __double_as_longlong(x)) ? y : x;
}

// Exact x/3 and x/6 (bit-identical to the IEEE quotient): Markstein's sequence
// q0 = RN(x*RN(1/d)), r = x - d*q0 (exact by FMA), q = RN(q0 + r*RN(1/d)); the
// sign is copied from x so that -0 is preserved.  Outside the exponent range
// where r is guaranteed exact (and for NaN/Inf) the true division is used.
__device__ __noinline__ double div_slow(double x, double d) { return x / d; }
template <int D>
__device__ __forceinline__ double divc(double x) {
  const double c = (D == 3) ? 0.33333333333333331 : 0.16666666666666666;
  const unsigned e = ((unsigned)__double2hiint(x) >> 20) & 0x7ffu;
  if (__builtin_expect((e - 100u) >= 1900u, 0)) {
    if (x == 0.0) return x;
    return div_slow(x, (double)D);
  }
  const double q0 = __dmul_rn(x, c);
  const double r = __fma_rn(-(double)D, q0, x);
  return copysign(__fma_rn(r, c, q0), x);
}

// PLM slope of a cell (:427-431 / :809-813) from the face differences
// a_hi = Tp-Tc, a_lo = Tc-Tm, their doubled magnitudes A2 = 2|a|, and d = Tp-Tm.
__device__ __forceinline__ double slope_f(double a_hi, double a_lo, double A2_hi, double A2_lo, double d, double mm) {
  double mag = 0.0;
  // same sign bit (a difference of two finite numbers is never -0)
  if ((__double2hiint(a_hi) ^ __double2hiint(a_lo)) >= 0) mag = minpos(minpos(0.5 * fabs(d), A2_hi), A2_lo);
  return mm * copysign(mag, d);
}

// CW84 edge value between cells lo and hi=lo+1 (:535-536 / :920-921)
__device__ __forceinline__ double edge_cw(double Tlo, double Thi, double slo, double shi) {
  return 0.5 * ((Tlo + Thi) + divc<3>(slo - shi));
}

// PPM:H3 edge values of one cell (:530-533 / :915-918)
__device__ __forceinline__ void edges_h3(double Tm, double Tc, double Tp, double &aL, double &aR) {
  aL = divc<6>(5. * Tc + (2. * Tm - Tp));
  aL = gmax2(gmin2(Tc, Tm), aL); aL = gmin2(gmax2(Tc, Tm), aL);
  aR = divc<6>(5. * Tc + (2. * Tp - Tm));
  aR = gmax2(gmin2(Tc, Tp), aR); aR = gmin2(gmax2(Tc, Tp), aR);
}

// Monotonic limiter (:538-548 / :923-933).  pp = mask*mask*(Tp-Tc)*(Tc-Tm), evaluated left to right.
// Returns the limited aL, aR, dA = aR-aL and a6.
__device__ __forceinline__ void ppm_limit(double &aL, double &aR, double Tc, double pp, double &dA, double &a6) {
  dA = aR - aL;
  double S = aR + aL;
  const double t1 = dA * (Tc - 0.5 * S), t2 = divc<6>(dA * dA);
  if (pp <= 0.) {            // PCM: aL = aR = Tc, so aR-aL = +0 and 6Tc - 3(Tc+Tc) = +0 (finite Tc)
    aL = Tc; aR = Tc; dA = 0.0; a6 = 0.0;
    return;
  }
  // |t1| > t2 (t2 >= +0): integer compare of the magnitudes
  if (__builtin_expect((__double_as_longlong(t1) & 0x7fffffffffffffffLL) > __double_as_longlong(t2), 0)) {
    if (__double2hiint(t1) >= 0) aL = 3. * Tc - 2. * aR;      // t1 > t2   (:543)
    else aR = 3. * Tc - 2. * aL;                              // t1 < -t2  (:545)
    dA = aR - aL; S = aR + aL;
  }
  a6 = 6. * Tc - 3. * S;
}

// Flux through a face whose upstream cell lies on its low side (hh >= 0, :551-552 / :936-938) ...
__device__ __forceinline__ double out_hi(double hh, double hc, double gq, double aR, double dA, double a6) {
  return hh * (aR - hc * (dA - a6 * gq));
}
// ... or on its high side (hh < 0, :554-555 / :940-942).  hc = 0.5*CFL, gq = 1 - 2/3*CFL.
__device__ __forceinline__ double out_lo(double hh, double hc, double gq, double aL, double dA, double a6) {
  return hh * (aL + hc * (dA + a6 * gq));
}

__device__ __forceinline__ double underflow(double o, double cu) {
  // abs(o) < cu with both magnitudes as integers (cu > 0)
  return ((__double_as_longlong(o) & 0x7fffffffffffffffLL) < __double_as_longlong(cu)) ? 0.0 : o;
}

// ---------------------------------------------------------------------------
// x sweep
// ---------------------------------------------------------------------------
#define XT_USE 120     // useful cells per 128-cell warp tile (lanes 1..30)

template <int SCHEME>
__global__ void __launch_bounds__(128, 4)
k_x_fast(Geom g, Range rg, TrTable tt, const double *__restrict__ hhv, const double *__restrict__ hcv,
         const double *__restrict__ gqv, const double *__restrict__ hlstv, const double *__restrict__ ihv,
         const double *__restrict__ mmv, const int *__restrict__ entry, const int *__restrict__ domore_k, int wbase) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int rj = blockIdx.y * 4 + (threadIdx.x >> 5), k0 = blockIdx.z;
  const int j = rj - g.roff;
  if (j < g.jsd || j > g.jed) return;                        // warp-uniform
  const int c0 = wbase + XT_USE * blockIdx.x + 4 * lane;       // first of this lane's 4 columns
  const long long q0 = (long long)c0 + (long long)g.P * rj + g.plane * k0;
  const bool lev = domore_k[k0] > 0;
  const bool in_j = (j >= rg.js && j <= rg.je);
  const bool active = lev && in_j && entry[k0 * g.R + rj];
  const bool st = (lane >= 1 && lane <= 30) && (c0 + 3 >= g.isd + g.coff) && (c0 <= g.ied + g.coff);
  bool inr[4];
#pragma unroll
  for (int n = 0; n < 4; n++) { const int i = c0 + n - g.coff; inr[n] = lev && in_j && i >= rg.is && i <= rg.ie; }

  if (!active) {     // copy through (+ underflow on the sweep's own range, :683-687)
    if (!st) return;
    for (int m = 0; m < tt.n; m++) {
      const double2 v0 = *reinterpret_cast<const double2 *>(tt.src[m] + q0);
      const double2 v1 = *reinterpret_cast<const double2 *>(tt.src[m] + q0 + 2);
      double o[4] = {v0.x, v0.y, v1.x, v1.y};
      const double cu = tt.cu[m];
      if (cu > 0.0) {
#pragma unroll
        for (int n = 0; n < 4; n++) if (inr[n]) o[n] = underflow(o[n], cu);
      }
      *reinterpret_cast<double2 *>(tt.dst[m] + q0) = make_double2(o[0], o[1]);
      *reinterpret_cast<double2 *>(tt.dst[m] + q0 + 2) = make_double2(o[2], o[3]);
    }
    return;
  }

  // tracer-independent quantities of this lane's 4 cells and their 5 faces
  double hh[5], hc[5], gq[5], hl[4], ih[4], mm[4];     // faces: [0] = west face of cell 0, [n+1] = east face of cell n
  {
    const double2 a0 = *reinterpret_cast<const double2 *>(hhv + q0), a1 = *reinterpret_cast<const double2 *>(hhv + q0 + 2);
    const double2 b0 = *reinterpret_cast<const double2 *>(hcv + q0), b1 = *reinterpret_cast<const double2 *>(hcv + q0 + 2);
    const double2 c0v = *reinterpret_cast<const double2 *>(gqv + q0), c1v = *reinterpret_cast<const double2 *>(gqv + q0 + 2);
    const double2 h0 = *reinterpret_cast<const double2 *>(hlstv + q0), h1 = *reinterpret_cast<const double2 *>(hlstv + q0 + 2);
    const double2 i0 = *reinterpret_cast<const double2 *>(ihv + q0), i1 = *reinterpret_cast<const double2 *>(ihv + q0 + 2);
    const long long q2 = (long long)c0 + (long long)g.P * rj;
    const double2 m0 = *reinterpret_cast<const double2 *>(mmv + q2), m1 = *reinterpret_cast<const double2 *>(mmv + q2 + 2);
    hh[1] = a0.x; hh[2] = a0.y; hh[3] = a1.x; hh[4] = a1.y;
    hc[1] = b0.x; hc[2] = b0.y; hc[3] = b1.x; hc[4] = b1.y;
    gq[1] = c0v.x; gq[2] = c0v.y; gq[3] = c1v.x; gq[4] = c1v.y;
    hl[0] = h0.x; hl[1] = h0.y; hl[2] = h1.x; hl[3] = h1.y;
    ih[0] = i0.x; ih[1] = i0.y; ih[2] = i1.x; ih[3] = i1.y;
    mm[0] = m0.x; mm[1] = m0.y; mm[2] = m1.x; mm[3] = m1.y;
    hh[0] = __shfl_up_sync(FULL, hh[4], 1);
    hc[0] = __shfl_up_sync(FULL, hc[4], 1);
    gq[0] = __shfl_up_sync(FULL, gq[4], 1);
  }

  double2 n0 = *reinterpret_cast<const double2 *>(tt.src[0] + q0);
  double2 n1 = *reinterpret_cast<const double2 *>(tt.src[0] + q0 + 2);
  for (int m = 0; m < tt.n; m++) {
    double T[6];      // T[0] = west neighbour, T[1..4] own cells, T[5] = east neighbour
    T[1] = n0.x; T[2] = n0.y; T[3] = n1.x; T[4] = n1.y;
    if (m + 1 < tt.n) {       // prefetch the next tracer
      n0 = *reinterpret_cast<const double2 *>(tt.src[m + 1] + q0);
      n1 = *reinterpret_cast<const double2 *>(tt.src[m + 1] + q0 + 2);
    }
    T[0] = __shfl_up_sync(FULL, T[4], 1);
    T[5] = __shfl_down_sync(FULL, T[1], 1);
    double a[5], A2[5];       // a[n] = T[n+1]-T[n]: difference across face n (same numbering as hh)
#pragma unroll
    for (int n = 0; n < 5; n++) { a[n] = T[n + 1] - T[n]; A2[n] = 2.0 * fabs(a[n]); }
    double s[4];
    if (SCHEME != SCHEME_PPM_H3) {
#pragma unroll
      for (int n = 0; n < 4; n++) s[n] = slope_f(a[n + 1], a[n], A2[n + 1], A2[n], T[n + 2] - T[n], mm[n]);
    }
    double fl[5];      // flux through face n
    if (SCHEME == SCHEME_PLM) {
      // face n is fed by cell n-1 (hh >= 0) or cell n (hh < 0); own cells are 0..3, cell 4 is the east neighbour's
      const double s_e = __shfl_down_sync(FULL, s[0], 1);
#pragma unroll
      for (int n = 1; n < 5; n++) {
        const double sup = (n < 4) ? s[n] : s_e, Tup = T[n + 1];
        fl[n] = (hh[n] >= 0.0) ? hh[n] * (T[n] + 0.5 * s[n - 1] * gq[n]) : hh[n] * (Tup - 0.5 * sup * gq[n]);
      }
    } else {
      double aL[4], aR[4], dA[4], a6[4];
      if (SCHEME == SCHEME_PPM_CW) {
        const double s_e = __shfl_down_sync(FULL, s[0], 1);
        double e[5];     // e[n+1] = edge value at the east face of cell n
#pragma unroll
        for (int n = 0; n < 4; n++) e[n + 1] = edge_cw(T[n + 1], T[n + 2], s[n], (n < 3) ? s[n + 1] : s_e);
        e[0] = __shfl_up_sync(FULL, e[4], 1);
#pragma unroll
        for (int n = 0; n < 4; n++) { aL[n] = e[n]; aR[n] = e[n + 1]; }
      } else {
#pragma unroll
        for (int n = 0; n < 4; n++) edges_h3(T[n], T[n + 1], T[n + 2], aL[n], aR[n]);
      }
#pragma unroll
      for (int n = 0; n < 4; n++) ppm_limit(aL[n], aR[n], T[n + 1], mm[n] * a[n + 1] * a[n], dA[n], a6[n]);
      // the east neighbour's first cell feeds face 4 when hh[4] < 0
      const double aL_e = __shfl_down_sync(FULL, aL[0], 1), dA_e = __shfl_down_sync(FULL, dA[0], 1),
                   a6_e = __shfl_down_sync(FULL, a6[0], 1);
#pragma unroll
      for (int n = 1; n < 5; n++) {
        const double xL = (n < 4) ? aL[n] : aL_e, xD = (n < 4) ? dA[n] : dA_e, x6 = (n < 4) ? a6[n] : a6_e;
        fl[n] = (hh[n] >= 0.0) ? out_hi(hh[n], hc[n], gq[n], aR[n - 1], dA[n - 1], a6[n - 1])
                               : out_lo(hh[n], hc[n], gq[n], xL, xD, x6);
      }
    }
    fl[0] = __shfl_up_sync(FULL, fl[4], 1);
    const double cu = tt.cu[m];
    double o[4];
#pragma unroll
    for (int n = 0; n < 4; n++) {
      o[n] = T[n + 1];
      if (ih[n] > 0.0) o[n] = (T[n + 1] * hl[n] - (fl[n + 1] - fl[n])) * ih[n];
      if (cu > 0.0 && inr[n]) o[n] = underflow(o[n], cu);
    }
    if (st) {
      *reinterpret_cast<double2 *>(tt.dst[m] + q0) = make_double2(o[0], o[1]);
      *reinterpret_cast<double2 *>(tt.dst[m] + q0 + 2) = make_double2(o[2], o[3]);
    }
  }
}

int x_fast_tiles(const Geom &g, int wbase) { return (g.ied + g.coff - (wbase + 4)) / XT_USE + 1; }

void launch_x_fast(cudaStream_t s, int scheme, const Geom &g, Range rg, const TrTable &tt, const double *hh,
                   const double *hc, const double *gq, const double *hlst, const double *ihnew, const double *mm,
                   const int *entry, const int *domore_k, int wbase) {
  dim3 grid(x_fast_tiles(g, wbase), (g.R + 3) / 4, g.nk);
  if (scheme == SCHEME_PLM) k_x_fast<SCHEME_PLM><<<grid, 128, 0, s>>>(g, rg, tt, hh, hc, gq, hlst, ihnew, mm, entry, domore_k, wbase);
  else if (scheme == SCHEME_PPM_CW) k_x_fast<SCHEME_PPM_CW><<<grid, 128, 0, s>>>(g, rg, tt, hh, hc, gq, hlst, ihnew, mm, entry, domore_k, wbase);
  else k_x_fast<SCHEME_PPM_H3><<<grid, 128, 0, s>>>(g, rg, tt, hh, hc, gq, hlst, ihnew, mm, entry, domore_k, wbase);
}

// ---------------------------------------------------------------------------
// y sweep: register pipeline marching north.  At step r (tracer row r in hand)
// the thread finishes the slope of cell r-1, the edge between cells r-2 and
// r-1, the parabola of cell r-2, the flux through face J = r-3 and the update
// of cell r-3.  Tracer rows are prefetched two steps ahead, mass rows one.
// ---------------------------------------------------------------------------
#define YW 8           // tracers (warps) per CTA

struct YMass { double2 hh, hc, gq, hl, ih, mm; };

template <int SCHEME>
__global__ void __launch_bounds__(32 * YW, 2)
k_y_fast(Geom g, Range rg, TrTable tt, const double *__restrict__ hhv, const double *__restrict__ hcv,
         const double *__restrict__ gqv, const double *__restrict__ hlstv, const double *__restrict__ ihv,
         const double *__restrict__ mmv, const int *__restrict__ entry, const int *__restrict__ domore_k) {
  const int k0 = blockIdx.z;
  if (domore_k[k0] <= 0) return;
  const int m = blockIdx.y * YW + threadIdx.y;
  if (m >= tt.n) return;
  const int c = 2 * (blockIdx.x * 32 + threadIdx.x);
  if (c + 1 >= g.P) return;
  const int i0 = c - g.coff;
  if (i0 + 1 < rg.is || i0 > rg.ie) return;         // nothing of this column pair is in the sweep
  const bool inr[2] = {i0 >= rg.is && i0 <= rg.ie, i0 + 1 >= rg.is && i0 + 1 <= rg.ie};
  double *__restrict__ Tt = tt.dst[m];              // in place
  const double cu = tt.cu[m];
  const int P = g.P;
  const long long base = (long long)c + g.plane * k0;      // + P*row
  const int *__restrict__ ent = entry + k0 * g.R;
  const int r_lo = rg.js - 3 + g.roff, r_hi = rg.je + 3 + g.roff;       // padded rows
  const int rf_lo = rg.js - 1 + g.roff, rf_hi = rg.je + g.roff;         // face rows of the sweep
  const int rc_lo = rg.js + g.roff;                                     // first cell row to update

  double T1[2] = {0, 0}, T2[2] = {0, 0}, T3[2] = {0, 0};   // T(r-1), T(r-2), T(r-3)
  double a1[2] = {0, 0}, A21[2] = {0, 0};                  // difference across face r-2|r-1 and 2|.|
  double s2[2] = {0, 0}, e3[2] = {0, 0}, pp2[2] = {0, 0};  // slope, south edge, limiter product of cell r-2
  double Fn3[2] = {0, 0}, flp[2] = {0, 0};                 // north outflow of cell r-3, flux through face r-4
  double hh3[2] = {0, 0}, hc3[2] = {0, 0}, gq3[2] = {0, 0};   // factors of face J = r-3
  bool act3 = false;                                       // face J = r-3 is part of the sweep

  double2 nx1 = *reinterpret_cast<const double2 *>(Tt + base + (long long)P * r_lo);
  double2 nx2 = *reinterpret_cast<const double2 *>(Tt + base + (long long)P * min(r_lo + 1, r_hi));
  double2 mmq = *reinterpret_cast<const double2 *>(mmv + c + (long long)P * max(r_lo - 1, 0));
#pragma unroll 3
  for (int r = r_lo; r <= r_hi; r++) {
    // loads first: they have the slope / edge / limiter arithmetic of this step to land
    const long long of = base + (long long)P * max(r - 2, 0), oc = base + (long long)P * max(r - 3, 0);
    YMass mc;
    mc.hh = *reinterpret_cast<const double2 *>(hhv + of);
    mc.hc = *reinterpret_cast<const double2 *>(hcv + of);
    mc.gq = *reinterpret_cast<const double2 *>(gqv + of);
    mc.hl = *reinterpret_cast<const double2 *>(hlstv + oc);
    mc.ih = *reinterpret_cast<const double2 *>(ihv + oc);
    mc.mm = mmq;
    mmq = *reinterpret_cast<const double2 *>(mmv + c + (long long)P * r);
    const double Tn[2] = {nx1.x, nx1.y};
    nx1 = nx2;
    nx2 = *reinterpret_cast<const double2 *>(Tt + base + (long long)P * min(r + 2, r_hi));
    const bool act2 = (r - 2 >= rf_lo && r - 2 <= rf_hi) && ent[r - 2];
    const bool out_row = (r - 3 >= rc_lo);
    const double hh2[2] = {mc.hh.x, mc.hh.y}, hc2[2] = {mc.hc.x, mc.hc.y}, gq2[2] = {mc.gq.x, mc.gq.y};
    const double hl[2] = {mc.hl.x, mc.hl.y}, ih[2] = {mc.ih.x, mc.ih.y}, mm1[2] = {mc.mm.x, mc.mm.y};
    double o[2];
#pragma unroll
    for (int n = 0; n < 2; n++) {
      const double a0 = Tn[n] - T1[n];                     // face r-1|r
      const double A20 = 2.0 * fabs(a0);
      double s1 = 0.0, Fn2, Fs2 = 0.0;
      if (SCHEME != SCHEME_PPM_H3) s1 = slope_f(a0, a1[n], A20, A21[n], Tn[n] - T2[n], mm1[n]);      // cell r-1
      if (SCHEME == SCHEME_PLM) {
        // cell r-2 feeds its north face r-2 (hh2 >= 0) and its south face r-3 (hh3 < 0); gq = 1-CFL
        Fn2 = hh2[n] * (T2[n] + 0.5 * s2[n] * gq2[n]);
        if (hh3[n] < 0.0) Fs2 = hh3[n] * (T2[n] - 0.5 * s2[n] * gq3[n]);
      } else {
        double aL, aR, dA, a6, e2 = 0.0;
        if (SCHEME == SCHEME_PPM_CW) { e2 = edge_cw(T2[n], T1[n], s2[n], s1); aL = e3[n]; aR = e2; }
        else edges_h3(T3[n], T2[n], T1[n], aL, aR);
        ppm_limit(aL, aR, T2[n], pp2[n], dA, a6);
        Fn2 = out_hi(hh2[n], hc2[n], gq2[n], aR, dA, a6);
        if (hh3[n] < 0.0) Fs2 = out_lo(hh3[n], hc3[n], gq3[n], aL, dA, a6);
        e3[n] = e2;
      }
      // flux through face J = r-3 (between cells r-3 and r-2); +0 on rows that are not swept (:1016-1018)
      const double fl3 = act3 ? ((hh3[n] >= 0.0) ? Fn3[n] : Fs2) : 0.0;
      o[n] = T3[n];
      if (ih[n] > 0.0) o[n] = (T3[n] * hl[n] - (fl3 - flp[n])) * ih[n];
      if (cu > 0.0 && inr[n]) o[n] = underflow(o[n], cu);
      flp[n] = fl3; Fn3[n] = Fn2;
      pp2[n] = mm1[n] * a0 * a1[n];                        // cell r-1: mask*(Tp-Tc)*(Tc-Tm)
      T3[n] = T2[n]; T2[n] = T1[n]; T1[n] = Tn[n];
      a1[n] = a0; A21[n] = A20; s2[n] = s1;
      hh3[n] = hh2[n]; hc3[n] = hc2[n]; gq3[n] = gq2[n];
    }
    act3 = act2;
    if (out_row) *reinterpret_cast<double2 *>(Tt + base + (long long)P * (r - 3)) = make_double2(o[0], o[1]);
  }
}

void launch_y_fast(cudaStream_t s, int scheme, const Geom &g, Range rg, const TrTable &tt, const double *hh,
                   const double *hc, const double *gq, const double *hlst, const double *ihnew, const double *mm,
                   const int *entry, const int *domore_k) {
  dim3 grid((g.P / 2 + 31) / 32, (tt.n + YW - 1) / YW, g.nk), block(32, YW);
  if (scheme == SCHEME_PLM) k_y_fast<SCHEME_PLM><<<grid, block, 0, s>>>(g, rg, tt, hh, hc, gq, hlst, ihnew, mm, entry, domore_k);
  else if (scheme == SCHEME_PPM_CW) k_y_fast<SCHEME_PPM_CW><<<grid, block, 0, s>>>(g, rg, tt, hh, hc, gq, hlst, ihnew, mm, entry, domore_k);
  else k_y_fast<SCHEME_PPM_H3><<<grid, block, 0, s>>>(g, rg, tt, hh, hc, gq, hlst, ihnew, mm, entry, domore_k);
}
