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

// min of two non-negative (sign bit clear) doubles
__device__ __forceinline__ double minpos(double x, double y) {
  return (__double_as_longlong(y) < __double_as_longlong(x)) ? y : x;
}

// PLM slope of a cell from a_hi = Tp-Tc, a_lo = Tc-Tm, d = Tp-Tm (:427-431 / :809-813)
__device__ __forceinline__ double slope_f(double a_hi, double a_lo, double d, double mm) {
  double mag = 0.0;
  // same sign bit (a difference of two numbers is never -0)
  if ((__double2hiint(a_hi) ^ __double2hiint(a_lo)) >= 0)
    mag = minpos(minpos(0.5 * fabs(d), 2.0 * fabs(a_hi)), 2.0 * fabs(a_lo));
  return mm * copysign(mag, d);
}

// CW84 edge value between cells lo and hi=lo+1 (:535-536 / :920-921)
__device__ __forceinline__ double edge_cw(double Tlo, double Thi, double slo, double shi) {
  return 0.5 * ((Tlo + Thi) + div3(slo - shi));
}

// PPM:H3 edge values of one cell (:530-533 / :915-918)
__device__ __forceinline__ void edges_h3(double Tm, double Tc, double Tp, double &aL, double &aR) {
  aL = div6(5. * Tc + (2. * Tm - Tp));
  aL = gmax2(gmin2(Tc, Tm), aL); aL = gmin2(gmax2(Tc, Tm), aL);
  aR = div6(5. * Tc + (2. * Tp - Tm));
  aR = gmax2(gmin2(Tc, Tp), aR); aR = gmin2(gmax2(Tc, Tp), aR);
}

// Monotonic limiter (:538-548 / :923-933): in aL, aR, Tc, a_hi = Tp-Tc, a_lo = Tc-Tm;
// out the limited aL, aR, dA = aR-aL and a6.
__device__ __forceinline__ void ppm_limit(double &aL, double &aR, double Tc, double a_hi, double a_lo, double mm,
                                          double &dA, double &a6) {
  dA = aR - aL;
  double S = aR + aL;
  const double mA = 0.5 * S;
  const double t1 = dA * (Tc - mA), t2 = div6(dA * dA);
  if (mm * a_hi * a_lo <= 0.) { aL = Tc; aR = Tc; dA = Tc - Tc; S = Tc + Tc; }
  else if (t1 > t2) { aL = 3. * Tc - 2. * aR; dA = aR - aL; S = aR + aL; }
  else if (t1 < -t2) { aR = 3. * Tc - 2. * aL; dA = aR - aL; S = aR + aL; }
  a6 = 6. * Tc - 3. * S;
}

// Outflow through the high (east / north) face of a cell, hh >= 0 (:551-552 / :936-938)
__device__ __forceinline__ double out_hi(double hh, double cfl, double aL, double aR, double dA, double a6) {
  const double c23 = 2. / 3.;
  (void)aL;
  return hh * (aR - 0.5 * cfl * (dA - a6 * (1. - c23 * cfl)));
}
// Outflow through the low (west / south) face of a cell, hh < 0 (:554-555 / :940-942)
__device__ __forceinline__ double out_lo(double hh, double cfl, double aL, double aR, double dA, double a6) {
  const double c23 = 2. / 3.;
  (void)aR;
  return hh * (aL + 0.5 * cfl * (dA + a6 * (1. - c23 * cfl)));
}

// ---------------------------------------------------------------------------
// x sweep
// ---------------------------------------------------------------------------
#define XT_USE 120     // useful cells per 128-cell warp tile (lanes 1..30)

template <int SCHEME>
__global__ void __launch_bounds__(128)
k_x_fast(Geom g, Range rg, TrTable tt, const double *__restrict__ hhv, const double *__restrict__ cflv,
         const double *__restrict__ hlstv, const double *__restrict__ ihv, const double *__restrict__ mmv,
         const int *__restrict__ entry, const int *__restrict__ domore_k, int wbase) {
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
    for (int m = 0; m < tt.n; m++) {
      if (!st) continue;
      const double2 v0 = *reinterpret_cast<const double2 *>(tt.src[m] + q0);
      const double2 v1 = *reinterpret_cast<const double2 *>(tt.src[m] + q0 + 2);
      double o[4] = {v0.x, v0.y, v1.x, v1.y};
      const double cu = tt.cu[m];
      if (cu > 0.0) {
#pragma unroll
        for (int n = 0; n < 4; n++) if (inr[n] && fabs(o[n]) < cu) o[n] = 0.0;
      }
      *reinterpret_cast<double2 *>(tt.dst[m] + q0) = make_double2(o[0], o[1]);
      *reinterpret_cast<double2 *>(tt.dst[m] + q0 + 2) = make_double2(o[2], o[3]);
    }
    return;
  }

  // tracer-independent quantities of this lane's 4 cells and their 5 faces
  double hh[5], cf[5], hl[4], ih[4], mm[4];     // faces: [0] = west face of cell 0, [n+1] = east face of cell n
  {
    const double2 a0 = *reinterpret_cast<const double2 *>(hhv + q0), a1 = *reinterpret_cast<const double2 *>(hhv + q0 + 2);
    const double2 b0 = *reinterpret_cast<const double2 *>(cflv + q0), b1 = *reinterpret_cast<const double2 *>(cflv + q0 + 2);
    const double2 h0 = *reinterpret_cast<const double2 *>(hlstv + q0), h1 = *reinterpret_cast<const double2 *>(hlstv + q0 + 2);
    const double2 i0 = *reinterpret_cast<const double2 *>(ihv + q0), i1 = *reinterpret_cast<const double2 *>(ihv + q0 + 2);
    const long long q2 = (long long)c0 + (long long)g.P * rj;
    const double2 m0 = *reinterpret_cast<const double2 *>(mmv + q2), m1 = *reinterpret_cast<const double2 *>(mmv + q2 + 2);
    hh[1] = a0.x; hh[2] = a0.y; hh[3] = a1.x; hh[4] = a1.y;
    cf[1] = b0.x; cf[2] = b0.y; cf[3] = b1.x; cf[4] = b1.y;
    hl[0] = h0.x; hl[1] = h0.y; hl[2] = h1.x; hl[3] = h1.y;
    ih[0] = i0.x; ih[1] = i0.y; ih[2] = i1.x; ih[3] = i1.y;
    mm[0] = m0.x; mm[1] = m0.y; mm[2] = m1.x; mm[3] = m1.y;
    hh[0] = __shfl_up_sync(FULL, hh[4], 1);
    cf[0] = __shfl_up_sync(FULL, cf[4], 1);
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
    double a[5];       // a[n] = T[n+1]-T[n]: difference across face n (same numbering as hh)
#pragma unroll
    for (int n = 0; n < 5; n++) a[n] = T[n + 1] - T[n];
    double aL[4], aR[4], dA[4], a6[4], s[4];
    if (SCHEME != SCHEME_PPM_H3) {
#pragma unroll
      for (int n = 0; n < 4; n++) s[n] = slope_f(a[n + 1], a[n], T[n + 2] - T[n], mm[n]);
    }
    if (SCHEME == SCHEME_PPM_CW) {
      const double s_e = __shfl_down_sync(FULL, s[0], 1);
      double e[5];     // e[n+1] = edge value at the east face of cell n
#pragma unroll
      for (int n = 0; n < 4; n++) e[n + 1] = edge_cw(T[n + 1], T[n + 2], s[n], (n < 3) ? s[n + 1] : s_e);
      e[0] = __shfl_up_sync(FULL, e[4], 1);
#pragma unroll
      for (int n = 0; n < 4; n++) { aL[n] = e[n]; aR[n] = e[n + 1]; }
    } else if (SCHEME == SCHEME_PPM_H3) {
#pragma unroll
      for (int n = 0; n < 4; n++) edges_h3(T[n], T[n + 1], T[n + 2], aL[n], aR[n]);
    }
    double Fe[4], Fw[4];     // outflow of cell n through its east / west face
    if (SCHEME == SCHEME_PLM) {
#pragma unroll
      for (int n = 0; n < 4; n++) {
        Fe[n] = hh[n + 1] * (T[n + 1] + 0.5 * s[n] * (1. - cf[n + 1]));
        Fw[n] = hh[n] * (T[n + 1] - 0.5 * s[n] * (1. - cf[n]));
      }
    } else {
#pragma unroll
      for (int n = 0; n < 4; n++) {
        ppm_limit(aL[n], aR[n], T[n + 1], a[n + 1], a[n], mm[n], dA[n], a6[n]);
        Fe[n] = out_hi(hh[n + 1], cf[n + 1], aL[n], aR[n], dA[n], a6[n]);
        Fw[n] = out_lo(hh[n], cf[n], aL[n], aR[n], dA[n], a6[n]);
      }
    }
    const double Fw_e = __shfl_down_sync(FULL, Fw[0], 1);      // west outflow of the east neighbour cell
    double fl[5];      // flux through face n
#pragma unroll
    for (int n = 1; n < 5; n++) fl[n] = (hh[n] >= 0.0) ? Fe[n - 1] : ((n < 4) ? Fw[n] : Fw_e);
    fl[0] = __shfl_up_sync(FULL, fl[4], 1);
    const double cu = tt.cu[m];
    double o[4];
#pragma unroll
    for (int n = 0; n < 4; n++) {
      o[n] = T[n + 1];
      if (ih[n] > 0.0) o[n] = (T[n + 1] * hl[n] - (fl[n + 1] - fl[n])) * ih[n];
      if (cu > 0.0 && inr[n] && fabs(o[n]) < cu) o[n] = 0.0;
    }
    if (st) {
      *reinterpret_cast<double2 *>(tt.dst[m] + q0) = make_double2(o[0], o[1]);
      *reinterpret_cast<double2 *>(tt.dst[m] + q0 + 2) = make_double2(o[2], o[3]);
    }
  }
}

int x_fast_tiles(const Geom &g, int wbase) { return (g.ied + g.coff - (wbase + 4)) / XT_USE + 1; }

void launch_x_fast(cudaStream_t s, int scheme, const Geom &g, Range rg, const TrTable &tt, const double *hh,
                   const double *cfl, const double *hlst, const double *ihnew, const double *mm, const int *entry,
                   const int *domore_k, int wbase) {
  dim3 grid(x_fast_tiles(g, wbase), (g.R + 3) / 4, g.nk);
  if (scheme == SCHEME_PLM) k_x_fast<SCHEME_PLM><<<grid, 128, 0, s>>>(g, rg, tt, hh, cfl, hlst, ihnew, mm, entry, domore_k, wbase);
  else if (scheme == SCHEME_PPM_CW) k_x_fast<SCHEME_PPM_CW><<<grid, 128, 0, s>>>(g, rg, tt, hh, cfl, hlst, ihnew, mm, entry, domore_k, wbase);
  else k_x_fast<SCHEME_PPM_H3><<<grid, 128, 0, s>>>(g, rg, tt, hh, cfl, hlst, ihnew, mm, entry, domore_k, wbase);
}

// ---------------------------------------------------------------------------
// y sweep: register pipeline marching north.  At step r (tracer row r just
// loaded) the thread finishes the slope of cell r-1, the edge between r-2 and
// r-1, the parabola and outflows of cell r-2, the flux through face J = r-3
// and the update of cell r-3.
// ---------------------------------------------------------------------------
#define YW 8           // tracers (warps) per CTA

template <int SCHEME>
__global__ void __launch_bounds__(32 * YW)
k_y_fast(Geom g, Range rg, TrTable tt, const double *__restrict__ hhv, const double *__restrict__ cflv,
         const double *__restrict__ hlstv, const double *__restrict__ ihv, const double *__restrict__ mmv,
         const int *__restrict__ entry, const int *__restrict__ domore_k) {
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
  const long long P = g.P;
  const long long base = (long long)c + g.plane * k0;      // + P*row
  const int *__restrict__ ent = entry + k0 * g.R;

  double T1[2] = {0, 0}, T2[2] = {0, 0}, T3[2] = {0, 0};   // T(r-1), T(r-2), T(r-3)
  double a1[2] = {0, 0}, a2[2] = {0, 0};                   // a(r-2|r-1), a(r-3|r-2)
  double s2[2] = {0, 0}, e3[2] = {0, 0};                   // slope of cell r-2, edge north of cell r-3
  double Fn3[2] = {0, 0}, flp[2] = {0, 0};                 // north outflow of cell r-3, flux through face r-4
  double hh3[2] = {0, 0}, cf3[2] = {0, 0};                 // vhh, CFL of face J = r-3
  double mm2[2] = {0, 0};                                  // mask product of cell r-2
  bool act3 = false;                                       // face J = r-3 is part of the sweep
  const int r_lo = rg.js - 3 + g.roff, r_hi = rg.je + 3 + g.roff;       // padded rows
  double2 nx = *reinterpret_cast<const double2 *>(Tt + base + P * r_lo);
  for (int r = r_lo; r <= r_hi; r++) {
    const double Tn[2] = {nx.x, nx.y};
    if (r < r_hi) nx = *reinterpret_cast<const double2 *>(Tt + base + P * (r + 1));
    // mass rows: face / cell r-2 (flux), cell r-3 (update), cell r-1 (slope mask)
    const int jf = r - 2 - g.roff;                         // Fortran index of face / cell r-2
    const bool act2 = (jf >= rg.js - 1 && jf <= rg.je) && ent[r - 2];
    double hh2[2] = {0, 0}, cf2[2] = {0, 0};
    if (act2) {
      const double2 u = *reinterpret_cast<const double2 *>(hhv + base + P * (r - 2));
      const double2 v = *reinterpret_cast<const double2 *>(cflv + base + P * (r - 2));
      hh2[0] = u.x; hh2[1] = u.y; cf2[0] = v.x; cf2[1] = v.y;
    }
    const double2 mq = *reinterpret_cast<const double2 *>(mmv + c + P * (r - 1));
    const double mm1[2] = {mq.x, mq.y};
    const int jc = r - 3 - g.roff;                         // cell to update
    const bool out_row = (jc >= rg.js && jc <= rg.je);
    double hl[2] = {0, 0}, ih[2] = {0, 0};
    if (out_row && (act2 || act3)) {     // (hh2, hh3 are the faces of cell r-3 ... its ih is 0 otherwise)
      const double2 u = *reinterpret_cast<const double2 *>(hlstv + base + P * (r - 3));
      const double2 v = *reinterpret_cast<const double2 *>(ihv + base + P * (r - 3));
      hl[0] = u.x; hl[1] = u.y; ih[0] = v.x; ih[1] = v.y;
    }
    double o[2];
#pragma unroll
    for (int n = 0; n < 2; n++) {
      const double a0 = Tn[n] - T1[n];                     // face r-1|r
      double aL, aR, dA = 0.0, a6 = 0.0, s1 = 0.0, Fn2, Fs2;
      if (SCHEME != SCHEME_PPM_H3) s1 = slope_f(a0, a1[n], Tn[n] - T2[n], mm1[n]);      // cell r-1
      if (SCHEME == SCHEME_PLM) {
        // cell r-2 with its own slope s2
        Fn2 = hh2[n] * (T2[n] + 0.5 * s2[n] * (1. - cf2[n]));
        Fs2 = hh3[n] * (T2[n] - 0.5 * s2[n] * (1. - cf3[n]));
      } else {
        double e2 = 0.0;
        if (SCHEME == SCHEME_PPM_CW) { e2 = edge_cw(T2[n], T1[n], s2[n], s1); aL = e3[n]; aR = e2; }
        else edges_h3(T3[n], T2[n], T1[n], aL, aR);
        ppm_limit(aL, aR, T2[n], a1[n], a2[n], mm2[n], dA, a6);
        Fn2 = out_hi(hh2[n], cf2[n], aL, aR, dA, a6);
        Fs2 = out_lo(hh3[n], cf3[n], aL, aR, dA, a6);
        e3[n] = e2;
      }
      // flux through face J = r-3 (between cells r-3 and r-2); +0 on rows that are not swept (:1016-1018)
      const double fl3 = act3 ? ((hh3[n] >= 0.0) ? Fn3[n] : Fs2) : 0.0;
      o[n] = T3[n];
      if (ih[n] > 0.0) o[n] = (T3[n] * hl[n] - (fl3 - flp[n])) * ih[n];
      if (cu > 0.0 && inr[n] && fabs(o[n]) < cu) o[n] = 0.0;
      flp[n] = fl3; Fn3[n] = Fn2;
      T3[n] = T2[n]; T2[n] = T1[n]; T1[n] = Tn[n];
      a2[n] = a1[n]; a1[n] = a0; s2[n] = s1;
      hh3[n] = hh2[n]; cf3[n] = cf2[n]; mm2[n] = mm1[n];
    }
    act3 = act2;
    if (out_row) *reinterpret_cast<double2 *>(Tt + base + P * (r - 3)) = make_double2(o[0], o[1]);
  }
}

void launch_y_fast(cudaStream_t s, int scheme, const Geom &g, Range rg, const TrTable &tt, const double *hh,
                   const double *cfl, const double *hlst, const double *ihnew, const double *mm, const int *entry,
                   const int *domore_k) {
  dim3 grid((g.P / 2 + 31) / 32, (tt.n + YW - 1) / YW, g.nk), block(32, YW);
  if (scheme == SCHEME_PLM) k_y_fast<SCHEME_PLM><<<grid, block, 0, s>>>(g, rg, tt, hh, cfl, hlst, ihnew, mm, entry, domore_k);
  else if (scheme == SCHEME_PPM_CW) k_y_fast<SCHEME_PPM_CW><<<grid, block, 0, s>>>(g, rg, tt, hh, cfl, hlst, ihnew, mm, entry, domore_k);
  else k_y_fast<SCHEME_PPM_H3><<<grid, block, 0, s>>>(g, rg, tt, hh, cfl, hlst, ihnew, mm, entry, domore_k);
}
