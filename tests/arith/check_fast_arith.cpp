// CPU check of the exact-by-construction rewrites used by the streaming kernels
// (mom6_b200/csrc/fast_common.cuh) against the literal restatement of the reference formulas
// (mom6_b200/csrc/internal.cuh: plm_slope / ppm_parabola / face_flux, which follow
// MOM_tracer_advect.F90:427-431,517-577).  Both headers are compiled as HOST code through a thin
// shim of the CUDA intrinsics they use, with -ffp-contract=off, so every double operation is the
// one the GPU executes under -fmad=false (explicit __fma_rn -> std::fma).
//
//   g++ -O2 -ffp-contract=off -D__CUDA_RUNTIME_H__ -I<repo>/mom6_b200/csrc check_fast_arith.cpp
//
// Test infrastructure only (the oracle is not involved: this pins the rewrites to the formulas
// they replace, bit for bit, on tens of millions of stencils incl. +-0, equal neighbours,
// denormals and masked cells).
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
typedef void *cudaStream_t;
#define __device__
#define __host__
#define __global__
#define __forceinline__ inline
#define __constant__
#define __restrict__
#define __align__(n)
static inline int __double2hiint(double x) { uint64_t b; memcpy(&b, &x, 8); return (int)(b >> 32); }
static inline int __double2loint(double x) { uint64_t b; memcpy(&b, &x, 8); return (int)(b & 0xffffffffu); }
static inline double __hiloint2double(int hi, int lo) {
  uint64_t b = ((uint64_t)(uint32_t)hi << 32) | (uint32_t)lo; double x; memcpy(&x, &b, 8); return x;
}
static inline long long __double_as_longlong(double x) { long long b; memcpy(&b, &x, 8); return b; }
static inline double __longlong_as_double(long long b) { double x; memcpy(&x, &b, 8); return x; }
static inline double __dmul_rn(double a, double b) { return a * b; }
static inline double __fma_rn(double a, double b, double c) { return std::fma(a, b, c); }
static inline double __shfl_up_sync(unsigned, double x, int) { return x; }      // (sweep4 is not instantiated here)
static inline double __shfl_down_sync(unsigned, double x, int) { return x; }
using std::copysign; using std::fabs;
#include "fast_common.cuh"

static uint64_t rs = 0x4D4F4D36ULL * 2654435761ULL + 12345;
static inline uint64_t rnd() { rs ^= rs << 13; rs ^= rs >> 7; rs ^= rs << 17; return rs; }
static inline double u01() { return (rnd() >> 11) * (1.0 / 9007199254740992.0); }
static inline bool same(double a, double b) { return memcmp(&a, &b, 8) == 0; }

// streaming formulation of the flux through the face between cells 3 and 4 of T[0..7]
// (cells 1..6 have slopes, cells 2..5 have edges, the upstream cell is 3 or 4), written with the
// helpers exactly as fast_x.cu / fast_y.cu use them.
template <int SCHEME>
static double stream_flux(const double *T, const double *mm, double hh, double cfl) {
  const double hc = 0.5 * cfl;
  const double gq = (SCHEME == SCHEME_PLM) ? 1. - cfl : 1. - (2. / 3.) * cfl;
  const bool lo = (hh >= 0.0);
  double a[8], A2[8];          // a[n] = T[n]-T[n-1]
  for (int n = 1; n < 8; n++) { a[n] = T[n] - T[n - 1]; A2[n] = 2.0 * fabs(a[n]); }
  double s[8] = {0};
  if (SCHEME != SCHEME_PPM_H3)
    for (int n = 1; n < 7; n++) s[n] = slope_f(a[n + 1], a[n], A2[n + 1], A2[n], T[n + 1] - T[n - 1], mm[n]);
  if (SCHEME == SCHEME_PLM) {
    const double gqs = neg_if(gq, !lo);
    const double sup = lo ? s[3] : s[4], Tup = lo ? T[3] : T[4];
    return hh * (Tup + 0.5 * sup * gqs);
  }
  double aL[8], aR[8], dA[8], a6[8];
  for (int c = 3; c <= 4; c++) {
    if (SCHEME == SCHEME_PPM_CW) {
      aL[c] = edge_cw(T[c - 1], T[c], s[c - 1], s[c]);
      aR[c] = edge_cw(T[c], T[c + 1], s[c], s[c + 1]);
    } else edges_h3(T[c - 1], T[c], T[c + 1], aL[c], aR[c]);
    double S, t1, dd, t2;
    ppm_pcm(aL[c], aR[c], T[c], mm[c] * a[c + 1] * a[c]);
    ppm_lim_a(aL[c], aR[c], T[c], dA[c], S, t1, dd, t2);
    if (div6_tiny_pos(dd)) t2 = div6_int(dd);
    if (ppm_over(t1, t2)) ppm_fix(aL[c], aR[c], T[c], t1, dA[c], S);
    a6[c] = ppm_a6(T[c], S);
  }
  return face_flux_sgn(hh, neg_if(hc, lo), neg_if(gq, lo), lo, aR[3], dA[3], a6[3], aL[4], dA[4], a6[4]);
}

template <int SCHEME>
static double literal_flux(const double *T, const double *mm, double hh, double cfl) {
  const double t[6] = {T[1], T[2], T[3], T[4], T[5], T[6]};
  // face_flux's mask arguments: SCHEME PPM reads mm[a-1], mm[a], mm[b], mm[b+1] with a = 3, b = 4;
  // ppm_parabola's outer slopes need the masks of cells 2..5
  const double m4[4] = {mm[2], mm[3], mm[4], mm[5]};
  return face_flux<SCHEME>(t, m4, hh, cfl);
}

static double pick_T(int mode, double base, double amp) {
  switch (mode) {
    case 0: return base + amp * (u01() - 0.5);                   // generic
    case 1: return base;                                         // flat
    case 2: return (rnd() & 1) ? base : base + amp;              // steps
    case 3: return (rnd() & 3) ? 0.0 : ((rnd() & 1) ? -0.0 : amp * u01());   // zeros of both signs
    case 4: return (u01() - 0.5) * 1e-308 * amp;                 // denormal range
    case 5: return (u01() - 0.5) * 4e-154;                       // products near the underflow threshold
    default: return base + amp * (u01() - 0.5) * 1e-12;          // nearly equal
  }
}

int main(int argc, char **argv) {
  const long ncase = argc > 1 ? atol(argv[1]) : 3000000;
  long bad = 0, n = 0;
  // 1. exact quotients
  for (long i = 0; i < ncase * 4; i++) {
    uint64_t b = rnd(), e;
    const int pass = i % 3;
    if (pass == 0) e = (b >> 52) & 0x7ff; else if (pass == 1) e = (b >> 52) & 0x3f; else e = 1000 + ((b >> 52) & 0x3f);
    if (e == 0x7ff) continue;
    uint64_t bits = (b & 0x800fffffffffffffULL) | (e << 52);
    if ((i & 1023) == 0) bits &= 0x8000000000000fffULL;          // deep denormals and +-0
    double x; memcpy(&x, &bits, 8);
    const double q3 = divc_fast<3>(x), q6 = divc<6>(x);
    if (!same(q3, x / 3.0) || !same(q6, x / 6.0)) { if (bad++ < 10) printf("div mismatch x=%a q3=%a q6=%a\n", x, q3, q6); }
    const double xx = fabs(x);                                     // a square is >= +0: the form the limiter uses
    double t2 = divc_fast<6>(xx); if (div6_tiny_pos(xx)) t2 = div6_int(xx);
    if (!same(t2, xx / 6.0)) { if (bad++ < 10) printf("div6 pos mismatch x=%a\n", xx); }
    n++;
  }
  // 2. slopes and fluxes
  for (long i = 0; i < ncase; i++) {
    double T[8], mm[8];
    const int mode = (int)(rnd() % 8);
    const double base = (rnd() & 1) ? 35.0 * u01() : (u01() - 0.5) * 60.0, amp = std::pow(10.0, -6.0 * u01()) * 10.0;
    for (int k = 0; k < 8; k++) {
      T[k] = pick_T((rnd() & 7) ? mode : (int)(rnd() % 8), base, amp);
      mm[k] = (rnd() % 5) ? 1.0 : 0.0;
    }
    const double cfl = (rnd() & 7) ? u01() : ((rnd() & 1) ? 0.0 : 1.0);
    double hh = (u01() - 0.5) * 1e6; if (!(rnd() & 15)) hh = (rnd() & 1) ? 0.0 : -0.0;
    for (int k = 1; k < 7; k++) {
      const double lit = plm_slope(T[k + 1], T[k], T[k - 1], mm[k]);
      const double a_hi = T[k + 1] - T[k], a_lo = T[k] - T[k - 1];
      const double str = slope_f(a_hi, a_lo, 2.0 * fabs(a_hi), 2.0 * fabs(a_lo), T[k + 1] - T[k - 1], mm[k]);
      if (!same(lit, str)) { if (bad++ < 10) printf("slope mismatch %a %a %a mm=%g: %a vs %a\n", T[k - 1], T[k], T[k + 1], mm[k], lit, str); }
    }
    const double f0 = literal_flux<SCHEME_PLM>(T, mm, hh, cfl), g0 = stream_flux<SCHEME_PLM>(T, mm, hh, cfl);
    const double f1 = literal_flux<SCHEME_PPM_CW>(T, mm, hh, cfl), g1 = stream_flux<SCHEME_PPM_CW>(T, mm, hh, cfl);
    const double f2 = literal_flux<SCHEME_PPM_H3>(T, mm, hh, cfl), g2 = stream_flux<SCHEME_PPM_H3>(T, mm, hh, cfl);
    if (!same(f0, g0)) { if (bad++ < 10) printf("PLM flux mismatch %a vs %a (hh=%a cfl=%a mode %d)\n", f0, g0, hh, cfl, mode); }
    if (!same(f1, g1)) { if (bad++ < 10) printf("PPM flux mismatch %a vs %a (hh=%a cfl=%a mode %d)\n", f1, g1, hh, cfl, mode); }
    if (!same(f2, g2)) { if (bad++ < 10) printf("PPM:H3 flux mismatch %a vs %a (hh=%a cfl=%a mode %d)\n", f2, g2, hh, cfl, mode); }
    n += 9;
  }
  printf("checked %ld values, %ld mismatches\n", n, bad);
  return bad ? 1 : 0;
}
