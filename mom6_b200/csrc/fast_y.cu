// y sweep of the streaming path (see fast_common.cuh for the design notes).
#include "fast_common.cuh"

// ---------------------------------------------------------------------------
// y sweep: register pipeline marching north, fed by a shared-memory row ring.
//
// CTA = 64 columns x YW tracers (one consumer warp per tracer, two columns per
// lane) + one producer warp.  For every step r the producer issues 512-byte
// bulk copies (cp.async.bulk -> UBLKCP) of tracer row r of each tracer and of
// the six tracer-independent rows the step needs (uhh/0.5CFL/1-2/3CFL of face
// r-2, hlst/Ihnew of cell r-3, mask product of cell r-1) into stage r % YNS,
// completing on that stage's `full` mbarrier; consumers release the stage on
// its `empty` mbarrier as soon as the values are in registers.  YNS rows are
// in flight per CTA, so HBM latency is hidden without prefetch registers, and
// the mass rows are fetched once per CTA instead of once per tracer.
//
// At step r a consumer thread finishes the slope of cell r-1, the edge between
// cells r-2 and r-1, the parabola of cell r-2, the flux through face J = r-3
// and the update of cell r-3 (written in place: rows <= r-3 are never read again).
// ---------------------------------------------------------------------------
#define YW 7            // tracers (consumer warps) per CTA: 7 + 1 producer warp = 256 threads
#define YNS 8           // ring stages
#define YCOLS 64        // columns per CTA
#define YROWB (YCOLS * 8)
#ifndef YUNROLL
#define YUNROLL 3
#endif
#define Y_PRAGMA_(x) _Pragma(#x)
#define Y_UNROLL(n) Y_PRAGMA_(unroll n)

struct __align__(16) YStage { double mass[6][YCOLS]; double T[YW][YCOLS]; };

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
k_y_ring(Geom g, Range rg, TrTable tt, const double *__restrict__ hhv, const double *__restrict__ hcv,
         const double *__restrict__ gqv, const double *__restrict__ hlstv, const double *__restrict__ ihv,
         const double *__restrict__ mmv, const int *__restrict__ entry, const int *__restrict__ domore_k) {
  extern __shared__ __align__(128) unsigned char y_smem[];
  YStage *stg = reinterpret_cast<YStage *>(y_smem);
  __shared__ __align__(8) unsigned long long bar_full[YNS], bar_empty[YNS];
  const int k0 = blockIdx.z;
  if (domore_k[k0] <= 0) return;                                  // CTA-uniform
  const int cbase = YCOLS * blockIdx.x;
  if (cbase + YCOLS - 1 - g.coff < rg.is || cbase - g.coff > rg.ie) return;   // no column of the tile is swept
  const int m0 = blockIdx.y * YW;
  const int nact = min(YW, tt.n - m0);
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < YNS; s++) { mbar_init(smem_u32(&bar_full[s]), 1); mbar_init(smem_u32(&bar_empty[s]), nact); }
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
  if (w >= nact) return;

  // ---- consumer warp: tracer m0+w, columns cbase + 2*lane, +1 ----
  const int m = m0 + w;
  const int c = cbase + 2 * lane;
  const int i0 = c - g.coff;
  const bool inr[2] = {i0 >= rg.is && i0 <= rg.ie, i0 + 1 >= rg.is && i0 + 1 <= rg.ie};
  double *__restrict__ Tt = tt.dst[m] + lev + 2 * lane;     // in place; + P*row
  const double cu = tt.cu[m];
  const int *__restrict__ ent = entry + k0 * g.R;
  const int rf_lo = rg.js - 1 + g.roff, rf_hi = rg.je + g.roff;         // face rows of the sweep
  const int rc_lo = rg.js + g.roff;                                     // first cell row to update

  double T1[2] = {0, 0}, T2[2] = {0, 0}, T3[2] = {0, 0};   // T(r-1), T(r-2), T(r-3)
  double a1[2] = {0, 0}, A21[2] = {0, 0};                  // difference across face r-2|r-1 and 2|.|
  double s2[2] = {0, 0}, e3[2] = {0, 0}, pp2[2] = {0, 0};  // slope, south edge, limiter product of cell r-2
  double Fn3[2] = {0, 0};                                  // north outflow of cell r-3 (through face r-3)
  double flp[2] = {0, 0};                                  // flux through face r-4
  double hh3[2] = {0, 0}, hc3[2] = {0, 0}, gq3[2] = {0, 0};   // factors of face J = r-3
  bool act3 = false;                                       // face J = r-3 is part of the sweep
  const bool do_cu = (cu > 0.0);
  Y_UNROLL(YUNROLL)
  for (int it = 0; it < nsteps; it++) {
    const int r = r_lo + it, s = it % YNS;
    mbar_wait(smem_u32(&bar_full[s]), (it / YNS) & 1);
    const YStage &q = stg[s];
    const double2 tn = *reinterpret_cast<const double2 *>(&q.T[w][2 * lane]);
    const double2 qhh = *reinterpret_cast<const double2 *>(&q.mass[0][2 * lane]);
    const double2 qhc = *reinterpret_cast<const double2 *>(&q.mass[1][2 * lane]);
    const double2 qgq = *reinterpret_cast<const double2 *>(&q.mass[2][2 * lane]);
    const double2 qhl = *reinterpret_cast<const double2 *>(&q.mass[3][2 * lane]);
    const double2 qih = *reinterpret_cast<const double2 *>(&q.mass[4][2 * lane]);
    const double2 qmm = *reinterpret_cast<const double2 *>(&q.mass[5][2 * lane]);
    __syncwarp();
    if (lane == 0) mbar_arrive(smem_u32(&bar_empty[s]));      // the stage may be refilled
    const bool act2 = (r - 2 >= rf_lo && r - 2 <= rf_hi) && ent[r - 2];
    const bool out_row = (r - 3 >= rc_lo);
    const double Tn[2] = {tn.x, tn.y};
    const double hh2[2] = {qhh.x, qhh.y}, hc2[2] = {qhc.x, qhc.y}, gq2[2] = {qgq.x, qgq.y};
    const double hl[2] = {qhl.x, qhl.y}, ih[2] = {qih.x, qih.y}, mm1[2] = {qmm.x, qmm.y};
    double a0[2], A20[2], s1[2] = {0, 0};
#pragma unroll
    for (int n = 0; n < 2; n++) {
      a0[n] = Tn[n] - T1[n];                               // face r-1|r
      A20[n] = 2.0 * fabs(a0[n]);
      if (SCHEME != SCHEME_PPM_H3) s1[n] = slope_f(a0[n], a1[n], A20[n], A21[n], Tn[n] - T2[n], mm1[n]);    // cell r-1
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
    double o[2];
#pragma unroll
    for (int n = 0; n < 2; n++) {
      // cell r-2 feeds its north face r-2 when hh2 >= 0 (kept for the next step) and its south face r-3 when hh3 < 0
      double Fn2, Fs2 = 0.0;
      if (SCHEME == SCHEME_PLM) {      // hh*(T +- 0.5*slope*(1-CFL)), gq = 1-CFL (:946-961)
        Fn2 = hh2[n] * (xR[n] + dA[n] * gq2[n]);
        if (hh3[n] < 0.0) Fs2 = hh3[n] * (xL[n] - dA[n] * gq3[n]);
      } else {
        Fn2 = out_hi(hh2[n], hc2[n], gq2[n], xR[n], dA[n], a6[n]);
        if (hh3[n] < 0.0) Fs2 = out_lo(hh3[n], hc3[n], gq3[n], xL[n], dA[n], a6[n]);
      }
      // flux through face J = r-3 (between cells r-3 and r-2); +0 on rows that are not swept (:1016-1018)
      const double f = (hh3[n] >= 0.0) ? Fn3[n] : Fs2;
      const double fl3 = act3 ? f : 0.0;
      const double upd = (T3[n] * hl[n] - (fl3 - flp[n])) * ih[n];
      o[n] = (__double_as_longlong(ih[n]) != 0LL) ? upd : T3[n];             // do_i: Ihnew > 0 (it is never negative)
      if (do_cu && inr[n]) o[n] = underflow(o[n], cu);
      flp[n] = fl3;
      Fn3[n] = Fn2;
      pp2[n] = mm1[n] * a0[n] * a1[n];                     // cell r-1: mask*(Tp-Tc)*(Tc-Tm)
      T3[n] = T2[n]; T2[n] = T1[n]; T1[n] = Tn[n];
      a1[n] = a0[n]; A21[n] = A20[n]; s2[n] = s1[n];
      hh3[n] = hh2[n]; hc3[n] = hc2[n]; gq3[n] = gq2[n];
    }
    act3 = act2;
    if (out_row) *reinterpret_cast<double2 *>(Tt + (long long)P * (r - 3)) = make_double2(o[0], o[1]);
  }
}

void launch_y_fast(cudaStream_t s, int scheme, const Geom &g, Range rg, const TrTable &tt, const double *hh,
                   const double *hc, const double *gq, const double *hlst, const double *ihnew, const double *mm,
                   const int *entry, const int *domore_k) {
  dim3 grid(g.P / YCOLS, (tt.n + YW - 1) / YW, g.nk), block(32 * (YW + 1));
  const size_t smem = sizeof(YStage) * YNS;
#define YLAUNCH(S)                                                                                        \
  do {                                                                                                    \
    static bool attr_set = false;                                                                         \
    if (!attr_set) { cudaFuncSetAttribute(k_y_ring<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); attr_set = true; } \
    k_y_ring<S><<<grid, block, smem, s>>>(g, rg, tt, hh, hc, gq, hlst, ihnew, mm, entry, domore_k);       \
  } while (0)
  if (scheme == SCHEME_PLM) YLAUNCH(SCHEME_PLM);
  else if (scheme == SCHEME_PPM_CW) YLAUNCH(SCHEME_PPM_CW);
  else YLAUNCH(SCHEME_PPM_H3);
#undef YLAUNCH
}
