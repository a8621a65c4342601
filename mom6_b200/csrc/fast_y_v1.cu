// y sweep of the streaming path (see fast_common_v1.cuh for the design notes).
#include "fast_common_v1.cuh"

// ---------------------------------------------------------------------------
// y sweep: register pipeline marching north, fed by a shared-memory row ring.
//
// CTA = 32 columns x 2*YW tracers (one consumer warp per PAIR of tracers, one
// column per lane) + one producer warp.  For every step r the producer issues
// 256-byte bulk copies (cp.async.bulk -> UBLKCP) of tracer row r of each tracer
// and of the six tracer-independent rows the step needs (uhh/0.5CFL/1-2/3CFL of
// face r-2, hlst/Ihnew of cell r-3, mask product of cell r-1) into stage
// r % YNS, completing on that stage's `full` mbarrier; consumers release the
// stage on its `empty` mbarrier as soon as the values are in registers.  YNS
// rows are in flight per CTA, so HBM latency is hidden without prefetch
// registers; the mass rows are fetched once per 14 tracers, and a thread holds
// them once for its two tracers (fewer registers and selects than two columns
// of one tracer).
//
// At step r a consumer thread finishes the slope of cell r-1, the edge between
// cells r-2 and r-1, the parabola of cell r-2, the flux through face J = r-3
// and the update of cell r-3 (written in place: rows <= r-3 are never read again).
// ---------------------------------------------------------------------------
#ifndef YW
#define YW 7            // consumer warps per CTA: 7 + 1 producer warp = 256 threads
#endif
#define YTW 2           // tracers per consumer warp
#define YT (YW * YTW)   // tracers per CTA
#define YNS 8           // ring stages
#define YCOLS 32        // columns per CTA
#define YROWB (YCOLS * 8)
#ifndef YUNROLL
#define YUNROLL 2
#endif
#define Y_PRAGMA_(x) _Pragma(#x)
#define Y_UNROLL(n) Y_PRAGMA_(unroll n)

struct __align__(16) YStage { double mass[6][YCOLS]; double T[YT][YCOLS]; };

__device__ __forceinline__ unsigned smem_u32(const void *p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned bar, unsigned parity) {
  unsigned ok;
  do {
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  } while (!ok);
}
__device__ __forceinline__ void mbar_wait_sleep(unsigned bar, unsigned parity) {      // producer: back off while the ring is full
  unsigned ok;
  for (;;) {
    asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                 : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (ok) break;
    __nanosleep(256);
  }
}
__device__ __forceinline__ void bulk_g2s(unsigned dst, const void *src, unsigned bytes, unsigned bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

#ifndef YMINB
#define YMINB 2
#endif
template <int SCHEME>
__global__ void __launch_bounds__(32 * (YW + 1), YMINB)
k_y_ring_v1(Geom g, Range rg, TrTable tt, const double *__restrict__ hhv, const double *__restrict__ hcv,
         const double *__restrict__ gqv, const double *__restrict__ hlstv, const double *__restrict__ ihv,
         const double *__restrict__ mmv, const int *__restrict__ entry, const int *__restrict__ domore_k) {
  extern __shared__ __align__(128) unsigned char y_smem[];
  YStage *stg = reinterpret_cast<YStage *>(y_smem);
  __shared__ __align__(8) unsigned long long bar_full[YNS], bar_empty[YNS];
  const int k0 = blockIdx.z;
  if (domore_k[k0] <= 0) return;                                  // CTA-uniform
  const int cbase = YCOLS * blockIdx.x;
  if (cbase + YCOLS - 1 - g.coff < rg.is || cbase - g.coff > rg.ie) return;   // no column of the tile is swept
  const int m0 = blockIdx.y * YT;
  const int nact = min(YT, tt.n - m0);                            // tracers of this CTA
  const int nwarp = (nact + YTW - 1) / YTW;                       // consumer warps with work
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < YNS; s++) { mbar_init(smem_u32(&bar_full[s]), 1); mbar_init(smem_u32(&bar_empty[s]), nwarp); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();
  const int P = g.P;
  const int r_lo = rg.js - 3 + g.roff, r_hi = rg.je + 3 + g.roff;       // padded rows of the march
  const int nsteps = r_hi - r_lo + 1;
  const long long lev = g.plane * k0 + cbase;

  if (w == YW) {
    // ---- producer warp: lane l < 6 owns mass row l, lane 6+t the tracer row of tracer m0+t ----
    const double *src = nullptr;
    int roff = 0;                   // row of the array this lane copies at step r is max(r - roff, 0)
    if (lane == 0) { src = hhv + lev; roff = 2; }
    else if (lane == 1) { src = hcv + lev; roff = 2; }
    else if (lane == 2) { src = gqv + lev; roff = 2; }
    else if (lane == 3) { src = hlstv + lev; roff = 3; }
    else if (lane == 4) { src = ihv + lev; roff = 3; }
    else if (lane == 5) { src = mmv + cbase; roff = 1; }
    else if (lane < 6 + nact) { src = tt.dst[m0 + lane - 6] + lev; roff = 0; }
    for (int it = 0; it < nsteps; it++) {
      const int s = it % YNS;
      mbar_wait_sleep(smem_u32(&bar_empty[s]), ((it / YNS) & 1) ^ 1);
      const unsigned fb = smem_u32(&bar_full[s]);
      if (lane == 0) mbar_expect_tx(fb, (6 + nact) * YROWB);
      __syncwarp();
      if (lane < 6 + nact) {
        const int row = max(r_lo + it - roff, 0);
        const unsigned dst = smem_u32(lane < 6 ? &stg[s].mass[lane][0] : &stg[s].T[lane - 6][0]);
        bulk_g2s(dst, src + (long long)P * row, YROWB, fb);
      }
    }
    return;
  }
  if (w >= nwarp) return;

  // ---- consumer warp: tracers m0+2w and m0+2w+1 (n = 0, 1), column cbase + lane ----
  const int t0 = YTW * w;
  const bool two = (t0 + 1 < nact);                          // the warp's second tracer exists
  const int ts[2] = {t0, two ? t0 + 1 : t0};                 // (a lone tracer is computed twice, stored once)
  const int i0 = cbase + lane - g.coff;
  const bool inr = (i0 >= rg.is && i0 <= rg.ie);
  double *__restrict__ Tt0 = tt.dst[m0 + ts[0]] + lev + lane;     // in place; + P*row
  double *__restrict__ Tt1 = tt.dst[m0 + ts[1]] + lev + lane;
  const double cu[2] = {tt.cu[m0 + ts[0]], tt.cu[m0 + ts[1]]};
  const bool do_cu[2] = {cu[0] > 0.0 && inr, cu[1] > 0.0 && inr};
  const int *__restrict__ ent = entry + k0 * g.R;
  const int rf_lo = rg.js - 1 + g.roff, rf_hi = rg.je + g.roff;         // face rows of the sweep
  const int rc_lo = rg.js + g.roff;                                     // first cell row to update

  double T1[2] = {0, 0}, T2[2] = {0, 0}, T3[2] = {0, 0};   // T(r-1), T(r-2), T(r-3)
  double a1[2] = {0, 0}, A21[2] = {0, 0};                  // difference across face r-2|r-1 and 2|.|
  double s2[2] = {0, 0}, e3[2] = {0, 0}, pp2[2] = {0, 0};  // slope, south edge, limiter product of cell r-2
  double Fn3[2] = {0, 0};                                  // north outflow of cell r-3 (through face r-3)
  double flp[2] = {0, 0};                                  // flux through face r-4
  double hh3 = 0, hc3 = 0, gq3 = 0;                        // factors of face J = r-3 (shared by both tracers)
  bool act3 = false;                                       // face J = r-3 is part of the sweep
  Y_UNROLL(YUNROLL)
  for (int it = 0; it < nsteps; it++) {
    const int r = r_lo + it, s = it % YNS;
    mbar_wait(smem_u32(&bar_full[s]), (it / YNS) & 1);
    const YStage &q = stg[s];
    const double Tn[2] = {q.T[ts[0]][lane], q.T[ts[1]][lane]};
    const double hh2 = q.mass[0][lane], hc2 = q.mass[1][lane], gq2 = q.mass[2][lane];
    const double hl = q.mass[3][lane], ih = q.mass[4][lane], mm1 = q.mass[5][lane];
    __syncwarp();
    if (lane == 0) mbar_arrive(smem_u32(&bar_empty[s]));      // the stage may be refilled
    const bool act2 = (r - 2 >= rf_lo && r - 2 <= rf_hi) && ent[r - 2];
    const bool out_row = (r - 3 >= rc_lo);
    double a0[2], A20[2], s1[2] = {0, 0};
#pragma unroll
    for (int n = 0; n < 2; n++) {
      a0[n] = Tn[n] - T1[n];                               // face r-1|r
      A20[n] = 2.0 * fabs(a0[n]);
      if (SCHEME != SCHEME_PPM_H3) s1[n] = slope_f(a0[n], a1[n], A20[n], A21[n], Tn[n] - T2[n], mm1);    // cell r-1
    }
    // parabola of cell r-2: xL/xR = edge values, dA, a6  (PLM: xL = xR = T, dA = 0.5*slope, a6 unused)
    double xL[2], xR[2], dA[2], a6[2] = {0, 0};
    if (SCHEME == SCHEME_PLM) {
#pragma unroll
      for (int n = 0; n < 2; n++) { xL[n] = T2[n]; xR[n] = T2[n]; dA[n] = 0.5 * s2[n]; }
    } else {
      double S[2], t1[2], dd[2], t2[2], e2[2] = {0, 0};
#pragma unroll
      for (int n = 0; n < 2; n++) {
        if (SCHEME == SCHEME_PPM_CW) { e2[n] = edge_cw(T2[n], T1[n], s2[n], s1[n]); xL[n] = e3[n]; xR[n] = e2[n]; }
        else edges_h3(T3[n], T2[n], T1[n], xL[n], xR[n]);
        ppm_lim_a(xL[n], xR[n], T2[n], dA[n], S[n], t1[n], dd[n], t2[n]);
      }
      if (div6_tiny(dd[0]) | div6_tiny(dd[1])) {           // denormal dA^2: rare
#pragma unroll
        for (int n = 0; n < 2; n++) if (div6_tiny(dd[n])) t2[n] = div6_int(dd[n]);
      }
      bool fix[2];
#pragma unroll
      for (int n = 0; n < 2; n++) fix[n] = ppm_lim_b(xL[n], xR[n], T2[n], pp2[n], dA[n], S[n], t1[n], t2[n]);
      if (fix[0] | fix[1]) {                               // overshoot correction: rare
#pragma unroll
        for (int n = 0; n < 2; n++) if (fix[n]) ppm_fix(xL[n], xR[n], T2[n], t1[n], dA[n], S[n]);
      }
#pragma unroll
      for (int n = 0; n < 2; n++) { a6[n] = ppm_a6(T2[n], S[n]); e3[n] = e2[n]; }
    }
    // cell r-2 feeds its north face r-2 when hh2 >= 0 (kept for the next step) and its south face r-3 when hh3 < 0
    const bool south = (hh3 < 0.0);                        // same for both tracers
    const bool do_i = (__double_as_longlong(ih) != 0LL);   // Ihnew > 0 (it is never negative)
    double o[2];
#pragma unroll
    for (int n = 0; n < 2; n++) {
      double Fn2, Fs2 = 0.0;
      if (SCHEME == SCHEME_PLM) {      // hh*(T +- 0.5*slope*(1-CFL)), gq = 1-CFL (:946-961)
        Fn2 = hh2 * (xR[n] + dA[n] * gq2);
        if (south) Fs2 = hh3 * (xL[n] - dA[n] * gq3);
      } else {
        Fn2 = out_hi(hh2, hc2, gq2, xR[n], dA[n], a6[n]);
        if (south) Fs2 = out_lo(hh3, hc3, gq3, xL[n], dA[n], a6[n]);
      }
      // flux through face J = r-3 (between cells r-3 and r-2); +0 on rows that are not swept (:1016-1018)
      const double f = south ? Fs2 : Fn3[n];
      const double fl3 = act3 ? f : 0.0;
      const double upd = (T3[n] * hl - (fl3 - flp[n])) * ih;
      o[n] = do_i ? upd : T3[n];
      if (do_cu[n]) o[n] = underflow(o[n], cu[n]);
      flp[n] = fl3;
      Fn3[n] = Fn2;
      pp2[n] = mm1 * a0[n] * a1[n];                        // cell r-1: mask*(Tp-Tc)*(Tc-Tm)
      T3[n] = T2[n]; T2[n] = T1[n]; T1[n] = Tn[n];
      a1[n] = a0[n]; A21[n] = A20[n]; s2[n] = s1[n];
    }
    hh3 = hh2; hc3 = hc2; gq3 = gq2;
    act3 = act2;
    if (out_row) {
      Tt0[(long long)P * (r - 3)] = o[0];
      if (two) Tt1[(long long)P * (r - 3)] = o[1];
    }
  }
}

void launch_y_fast_v1(cudaStream_t s, int scheme, const Geom &g, Range rg, const TrTable &tt, const double *hh,
                   const double *hc, const double *gq, const double *hlst, const double *ihnew, const double *mm,
                   const int *entry, const int *domore_k) {
  dim3 grid(g.P / YCOLS, (tt.n + YT - 1) / YT, g.nk), block(32 * (YW + 1));
  const size_t smem = sizeof(YStage) * YNS;
#define YLAUNCH(S)                                                                                        \
  do {                                                                                                    \
    static bool attr_set = false;                                                                         \
    if (!attr_set) { cudaFuncSetAttribute(k_y_ring_v1<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); attr_set = true; } \
    k_y_ring_v1<S><<<grid, block, smem, s>>>(g, rg, tt, hh, hc, gq, hlst, ihnew, mm, entry, domore_k);       \
  } while (0)
  if (scheme == SCHEME_PLM) YLAUNCH(SCHEME_PLM);
  else if (scheme == SCHEME_PPM_CW) YLAUNCH(SCHEME_PPM_CW);
  else YLAUNCH(SCHEME_PPM_H3);
#undef YLAUNCH
}
