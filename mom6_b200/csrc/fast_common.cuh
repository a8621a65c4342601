// Streaming tracer-sweep kernels (the production path when no flux diagnostic
// is associated).  Same arithmetic as kernels.cu / the reference
// (/root/reference/src/tracer/MOM_tracer_advect.F90:517-577,652-662 for x,
// :902-963,1043-1047 for y), restructured so that every slope, edge value,
// parabola and face flux is computed exactly ONCE per cell / face:
//
//   k_x_fast  (fast_x.cu) one warp = 128 consecutive cells of one row (4 per lane, LDG.128),
//             all tracers looped in-thread with the tracer-independent mass
//             quantities (uhh, signed 0.5CFL and 1-2/3CFL, hlst, Ihnew, mask product)
//             and one register of per-cell / per-face flags held in registers;
//             the per-tracer arithmetic is sweep4 below: neighbour slopes / edges /
//             outflows cross lanes by warp shuffle.  Out of place (buf[src] ->
//             buf[1-src]); tiles overlap by 8 columns (120 useful of 128).
//   k_y_ring  (fast_y.cu) one thread = one column x two tracers, marching in j with the
//             whole stencil pipeline in registers, fed by a shared-memory ring of rows
//             that a producer warp fills with bulk copies; in place.
//
// The cost model these kernels were shaped by: an FP64 instruction holds its scheduler's
// issue port for two cycles, so a sweep costs (non-FP64 instructions) + 2 x (FP64 instructions); the FP64 count
// is fixed by the reference's formulas, the rest is what the rewrites below remove.
//
// Exact-by-construction rewrites used (each is checked bit for bit against the literal
// formulas of internal.cuh on the CPU by tests/arith/check_fast_arith.cpp):
//   * max(Tp,Tc,Tm)-Tc and Tc-min(Tp,Tc,Tm) are obtained from the face
//     differences a_hi = Tp-Tc and a_lo = Tc-Tm (rounding is monotone and
//     sign-symmetric), so the PLM slope magnitude is
//     min(0.5|Tp-Tm|, 2|a_hi|, 2|a_lo|) when a_hi and a_lo have the same sign
//     and 0 otherwise; its sign is overwritten by sign(.., Tp-Tm) in the
//     reference, so zero signs cannot differ.
//   * min over non-negative doubles is an integer compare of the bit patterns.
//   * the CW84 edge value aR(i) and aL(i+1) are the same expression.
//   * a face's flux is evaluated by its upstream cell only, with the factors signed once per face.
//   * PCM is applied to the edge values before the limiter; the slope is branch-free; /3 and /6 are exact
//     without copysign (see the helpers).
#pragma once
#include "internal.cuh"
#include "kernels.h"

// min of two non-negative (sign bit clear) doubles: integer compare of the bit patterns
__device__ __forceinline__ double minpos(double x, double y) {
  return (__double_as_longlong(y) < __double_as_longlong(x)) ? y : x;
}

// Exact x/3 and x/6 (bit-identical to the IEEE quotient): Markstein's sequence
// q0 = RN(x*RN(1/d)), r = x - d*q0 (exact by FMA), q = RN(q0 + r*RN(1/d)); the
// sign is copied from x so that -0 is preserved.  Checked against true division
// on 6e8 random values over every exponent incl. denormals
// (scripts/check_markstein.c): d = 3 never differs; d = 6 differs only on exact
// ties of denormal quotients, |x| < 2^-1021.  There the bit pattern of |x| IS
// the integer |x| / 2^-1074, so the quotient is an integer division rounded
// half to even (div6_int) -- no call, no FP division in the hot loops.
static __constant__ double k_recip[2] = {0.33333333333333331, 0.16666666666666666};
template <int D>
__device__ __forceinline__ double divc_fast(double x) {
  // The remainder is formed with the opposite sign, rn = d*q0 - x (exact), and subtracted: q = q0 - rn*c.
  // Same value as q0 + r*c, and the zero cases come out with the sign of x without a copysign:
  // x = -0: q0 = -0, rn = +0, q = -0 - (+0)*c = -0;  x = +0: q = +0 - (+0)*c = +0.
  // RN(1/3), RN(1/6) come from the constant bank: an FP64 instruction takes a c[][] operand directly, whereas a
  // literal whose low word is not zero has to be rebuilt in a register pair (two moves per use under register pressure)
  const double c = (D == 3) ? k_recip[0] : k_recip[1];
  const double q0 = __dmul_rn(x, c);
  const double rn = __fma_rn((double)D, q0, -x);
  return __fma_rn(-rn, c, q0);
}
__device__ __forceinline__ bool div6_tiny(double x) {        // 0 < |x| < 2^-1021
  const unsigned u = (unsigned)__double2hiint(x) & 0x7fffffffu;
  return u < 0x00200000u && (u | (unsigned)__double2loint(x)) != 0u;
}
__device__ __forceinline__ bool div6_tiny_pos(double x) {    // same for x >= +0 (a square): no sign mask
  const unsigned u = (unsigned)__double2hiint(x);
  return u < 0x00200000u && (u | (unsigned)__double2loint(x)) != 0u;
}
__device__ __forceinline__ double div6_int(double x) {       // exact x/6 for |x| < 2^-1021
  const unsigned long long b = (unsigned long long)__double_as_longlong(x);
  const unsigned long long k = b & 0x7fffffffffffffffULL;
  unsigned long long q = k / 6ULL;
  const unsigned long long rem = k - 6ULL * q;
  q += (rem > 3ULL || (rem == 3ULL && (q & 1ULL))) ? 1ULL : 0ULL;
  return __longlong_as_double((long long)((b & 0x8000000000000000ULL) | q));
}
template <int D>
__device__ __forceinline__ double divc(double x) {           // general form (cold paths, PPM:H3 edges)
  double q = divc_fast<D>(x);
  if (D == 6 && div6_tiny(x)) q = div6_int(x);
  return q;
}

// PLM slope of a cell (:427-431 / :809-813) from the face differences
// a_hi = Tp-Tc, a_lo = Tc-Tm, their doubled magnitudes A2 = 2|a|, and d = Tp-Tm.
// Branch-free: when a_hi and a_lo differ in sign the candidate 0.5|d| is formed as |d|*0 = +0,
// whose bit pattern is the integer minimum, so the magnitude comes out +0 with no select.
__device__ __forceinline__ double slope_f(double a_hi, double a_lo, double A2_hi, double A2_lo, double d, double mm) {
  // same sign bit (a difference of two finite numbers is never -0)
  const int same = ((__double2hiint(a_hi) ^ __double2hiint(a_lo)) >= 0) ? 0x3fe00000 : 0;
  const double hd = fabs(d) * __hiloint2double(same, 0);
  const double mag = minpos(minpos(hd, A2_hi), A2_lo);
  return mm * copysign(mag, d);
}

// CW84 edge value between cells lo and hi=lo+1 (:535-536 / :920-921)
__device__ __forceinline__ double edge_cw(double Tlo, double Thi, double slo, double shi) {
  return 0.5 * ((Tlo + Thi) + divc_fast<3>(slo - shi));
}

// PPM:H3 edge values of one cell (:530-533 / :915-918)
__device__ __forceinline__ void edges_h3(double Tm, double Tc, double Tp, double &aL, double &aR) {
  aL = divc<6>(5. * Tc + (2. * Tm - Tp));
  aL = gmax2(gmin2(Tc, Tm), aL); aL = gmin2(gmax2(Tc, Tm), aL);
  aR = divc<6>(5. * Tc + (2. * Tp - Tm));
  aR = gmax2(gmin2(Tc, Tp), aR); aR = gmin2(gmax2(Tc, Tp), aR);
}

// Monotonic limiter (:538-548 / :923-933).  The PCM case (mask*(Tp-Tc)*(Tc-Tm) <= 0) is applied to
// the edge values FIRST: with aL = aR = Tc the reference's post-limiter quantities follow from the
// same arithmetic (aR-aL = +0, aR+aL = Tc+Tc, 6Tc-3(Tc+Tc) = +0) and the overshoot tests see
// t1 = +0*(Tc-Tc) = +0 and t2 = +0, i.e. neither fires, exactly as the elseif chain demands.
//   ppm_pcm   : aL, aR := Tc where pp <= 0
//   ppm_lim_a : dA, S, t1 = dA*(Tc - 0.5*S), dd = dA*dA, t2 = dd/6 (fast form)
//   [if div6_tiny_pos(dd): t2 = div6_int(dd)]      (rare: one branch per thread-step)
//   ppm_over  : |t1| > t2
//   [if flag: ppm_fix]                              (rare)
//   ppm_a6
__device__ __forceinline__ void ppm_pcm(double &aL, double &aR, double Tc, double pp) {
  const bool pcm = (pp <= 0.);
  aL = pcm ? Tc : aL; aR = pcm ? Tc : aR;
}
__device__ __forceinline__ void ppm_lim_a(double aL, double aR, double Tc, double &dA, double &S, double &t1,
                                          double &dd, double &t2) {
  dA = aR - aL;
  S = aR + aL;
  t1 = dA * (Tc - 0.5 * S);
  dd = dA * dA;
  t2 = divc_fast<6>(dd);
}
__device__ __forceinline__ bool ppm_over(double t1, double t2) { return fabs(t1) > t2; }
// where the flag is set: t1 > t2 -> aL = 3Tc - 2aR (:543); t1 < -t2 -> aR = 3Tc - 2aL (:545)
__device__ __forceinline__ void ppm_fix(double &aL, double &aR, double Tc, double t1, double &dA, double &S) {
  if (__double2hiint(t1) >= 0) aL = 3. * Tc - 2. * aR;
  else aR = 3. * Tc - 2. * aL;
  dA = aR - aL; S = aR + aL;
}
__device__ __forceinline__ double ppm_a6(double Tc, double S) { return 6. * Tc - 3. * S; }   // :548 / :933

// Flux through one face (:550-556 / :935-943) from the parabola of its upstream
// cell: the low-side cell (aR, dA, a6) when hh >= 0, else the high-side cell
// (aL, dA, a6).  hc = 0.5*CFL, gq = 1 - 2/3*CFL.  x - y == x + (-y) and (-x)*y == -(x*y)
// exactly, so with the tracer-independent factors signed once per face,
//   hcs = hh >= 0 ? -hc : hc,   gqs = hh >= 0 ? -gq : gq,
// both forms are the one expression hh*(X + hcs*(D + A6*gqs)):
//   hh >= 0:  hh*(aR - hc*(dA - a6*gq))      hh < 0:  hh*(aL + hc*(dA + a6*gq))
__device__ __forceinline__ double neg_if(double x, bool p) {
  return __hiloint2double(__double2hiint(x) ^ (p ? (int)0x80000000 : 0), __double2loint(x));
}
__device__ __forceinline__ double face_flux_sgn(double hh, double hcs, double gqs, bool up_lo, double aR_lo,
                                                double dA_lo, double a6_lo, double aL_hi, double dA_hi, double a6_hi) {
  const double X = up_lo ? aR_lo : aL_hi, D = up_lo ? dA_lo : dA_hi, A6 = up_lo ? a6_lo : a6_hi;
  return hh * (X + hcs * (D + A6 * gqs));
}
// the two one-sided forms (y march, where the two sides are evaluated on different steps)
__device__ __forceinline__ double out_hi(double hh, double hc, double gq, double aR, double dA, double a6) {
  return hh * (aR - hc * (dA - a6 * gq));
}
__device__ __forceinline__ double out_lo(double hh, double hc, double gq, double aL, double dA, double a6) {
  return hh * (aL + hc * (dA + a6 * gq));
}

__device__ __forceinline__ double underflow(double o, double cu) {
  // abs(o) < cu with both magnitudes as integers (cu > 0)
  return ((__double_as_longlong(o) & 0x7fffffffffffffffLL) < __double_as_longlong(cu)) ? 0.0 : o;
}

// ---------------------------------------------------------------------------
// One tracer, the 4 consecutive cells of this lane along the sweep direction (:411-433,517-577,652-662 for x,
// :794-822,902-963,1021-1047 for y), shared by the x sweep (cells = 4 columns) and the transposed y sweep (cells =
// 4 rows).  The 32 lanes of the warp hold 128 consecutive cells; neighbour values cross lanes by shuffle, so all
// 32 lanes must call together.  T[1..4] in; T[0], T[5] are fetched here.  Faces: [0] = low face of cell 0,
// [n+1] = high face of cell n.  flg: bit n (0..4) face n fed by its low-side cell; bit 5+n cell n updated;
// bit 13+n (FACEACT only) face n belongs to a swept row, else its flux is +0 (:1016-1018).
// ---------------------------------------------------------------------------
template <int SCHEME, bool FACEACT>
__device__ __forceinline__ void sweep4(double (&T)[6], const double (&hh)[5], const double (&hcs)[5], const double (&gqs)[5],
                                       const double (&hl)[4], const double (&ih)[4], const double (&mm)[4],
                                       const unsigned flg, double (&o)[4], double (&fl)[5]) {
  const unsigned FULL = 0xffffffffu;
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
    // fl[n]: flux through face n (returned: the flux diagnostics of the DIAG sweeps use it)
    if (SCHEME == SCHEME_PLM) {
      // face n is fed by cell n-1 (hh >= 0) or cell n (hh < 0); own cells are 0..3, cell 4 is the east neighbour's
      const double s_e = __shfl_down_sync(FULL, s[0], 1);
#pragma unroll
      for (int n = 1; n < 5; n++) {
        const bool lo = (flg >> n) & 1u;
        const double sup = lo ? s[n - 1] : ((n < 4) ? s[n] : s_e), Tup = lo ? T[n] : T[n + 1];
        fl[n] = hh[n] * (Tup + 0.5 * sup * gqs[n]);
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
      double S[4], t1[4], dd[4], t2[4];
#pragma unroll
      for (int n = 0; n < 4; n++) {
        ppm_pcm(aL[n], aR[n], T[n + 1], mm[n] * a[n + 1] * a[n]);
        ppm_lim_a(aL[n], aR[n], T[n + 1], dA[n], S[n], t1[n], dd[n], t2[n]);
      }
      if (div6_tiny_pos(dd[0]) | div6_tiny_pos(dd[1]) | div6_tiny_pos(dd[2]) | div6_tiny_pos(dd[3])) {   // denormal dA^2: rare
#pragma unroll
        for (int n = 0; n < 4; n++) if (div6_tiny_pos(dd[n])) t2[n] = div6_int(dd[n]);
      }
      bool anyfix = ppm_over(t1[0], t2[0]);
#pragma unroll
      for (int n = 1; n < 4; n++) anyfix |= ppm_over(t1[n], t2[n]);
      if (anyfix) {       // overshoot corrections are rare: one region per 4 cells
#pragma unroll
        for (int n = 0; n < 4; n++) {
          double u = t1[n];
          asm volatile("" : "+d"(u));      // re-test here instead of carrying four predicates through the hot path
          if (ppm_over(u, t2[n])) ppm_fix(aL[n], aR[n], T[n + 1], u, dA[n], S[n]);
        }
      }
#pragma unroll
      for (int n = 0; n < 4; n++) a6[n] = ppm_a6(T[n + 1], S[n]);
      // the east neighbour's first cell feeds face 4 when hh[4] < 0
      const double aL_e = __shfl_down_sync(FULL, aL[0], 1), dA_e = __shfl_down_sync(FULL, dA[0], 1),
                   a6_e = __shfl_down_sync(FULL, a6[0], 1);
#pragma unroll
      for (int n = 1; n < 5; n++) {
        const double xL = (n < 4) ? aL[n] : aL_e, xD = (n < 4) ? dA[n] : dA_e, x6 = (n < 4) ? a6[n] : a6_e;
        fl[n] = face_flux_sgn(hh[n], hcs[n], gqs[n], (flg >> n) & 1u, aR[n - 1], dA[n - 1], a6[n - 1], xL, xD, x6);
      }
    }
    if (FACEACT) {
#pragma unroll
      for (int n = 1; n < 5; n++) fl[n] = ((flg >> (13 + n)) & 1u) ? fl[n] : 0.0;
    }
    fl[0] = __shfl_up_sync(FULL, fl[4], 1);
#pragma unroll
    for (int n = 0; n < 4; n++) {
      const double upd = (T[n + 1] * hl[n] - (fl[n + 1] - fl[n])) * ih[n];
      o[n] = ((flg >> (5 + n)) & 1u) ? upd : T[n + 1];
    }
}

template <int SCHEME, bool FACEACT>
__device__ __forceinline__ void sweep4(double (&T)[6], const double (&hh)[5], const double (&hcs)[5], const double (&gqs)[5],
                                       const double (&hl)[4], const double (&ih)[4], const double (&mm)[4],
                                       const unsigned flg, double (&o)[4]) {
  double fl[5];
  sweep4<SCHEME, FACEACT>(T, hh, hcs, gqs, hl, ih, mm, flg, o, fl);
}
