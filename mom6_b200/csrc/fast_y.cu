// y sweep of the streaming path (see fast_common.cuh for the design notes).
#include "fast_common.cuh"

// ---------------------------------------------------------------------------
// y sweep: register pipeline marching north, fed by a shared-memory row ring.
//
// CTA = 32 columns x 2*YW tracers (one consumer warp per PAIR of tracers, one
// column per lane) + one producer warp.  For every step r the producer issues
// (ONE elected lane, addresses in uniform registers: no per-lane serialisation in front of UBLKCP)
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
// Consumers address the ring by byte offset (ld.shared from a stage base), read the per-step "face row is swept"
// flag from a shared-memory byte filled once per CTA, and keep conc_underflow out of line.  Measured and dropped:
// two rows per stage, 12 stages, 6 consumer warps, an early poll of the next stage, a row-window skip of inactive
// stretches, a transposed (sweep4 on columns) kernel -- profiles/r01_final_ncu.md.
// ---------------------------------------------------------------------------
#ifndef YW
#define YW 7            // consumer warps per CTA: 7 + 1 producer warp = 256 threads
#endif
#ifndef YTW
#define YTW 2           // tracers per consumer warp
#endif
#define YT (YW * YTW)   // tracers per CTA
#ifndef YNS
#define YNS 8           // ring stages
#endif
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

__device__ __forceinline__ bool elect_one() {
  unsigned p;
  asm volatile("{\n .reg .pred q;\n elect.sync _|q, 0xffffffff;\n selp.u32 %0, 1, 0, q;\n}" : "=r"(p));
  return p != 0;
}
__device__ __forceinline__ double lds_f64(unsigned addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr) : "memory");
  return v;
}

// conc_underflow test kept out of line: tracers with a threshold are the exception, and an inlined test is
// if-converted into ~10 predicated instructions that every step of every tracer would issue
__device__ __noinline__ double underflow_rare(double o, double cu, bool inr) { return inr ? underflow(o, cu) : o; }

#ifndef YMINB
#define YMINB 2
#endif
template <int SCHEME>
__global__ void __launch_bounds__(32 * (YW + 1), YMINB)
k_y_ring(Geom g, Range rg, TrTable tt, const double *__restrict__ hhv, const double *__restrict__ hcv,
         const double *__restrict__ gqv, const double *__restrict__ hlstv, const double *__restrict__ ihv,
         const double *__restrict__ mmv, const int *__restrict__ entry, const int *__restrict__ domore_k) {
  extern __shared__ __align__(128) unsigned char y_smem[];
  YStage *stg = reinterpret_cast<YStage *>(y_smem);      // (used by the producer; consumers address the ring by offset)
  __shared__ __align__(8) unsigned long long bar_full[YNS], bar_empty[YNS];
  const int k0 = blockIdx.z;
  if (domore_k[k0] <= 0) return;                                  // CTA-uniform
#ifdef YSWAP
  // tracer groups of one column tile are adjacent in launch order: they march together and share the mass rows in L2
  const int cbase = YCOLS * blockIdx.y;
#else
  const int cbase = YCOLS * blockIdx.x;
#endif
  if (cbase + YCOLS - 1 - g.coff < rg.is || cbase - g.coff > rg.ie) return;   // no column of the tile is swept
#ifdef YSWAP
  const int m0 = blockIdx.x * YT;
#else
  const int m0 = blockIdx.y * YT;
#endif
  const int nact = min(YT, tt.n - m0);                            // tracers of this CTA
  const int nwarp = (nact + YTW - 1) / YTW;                       // consumer warps with work
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < YNS; s++) { mbar_init(smem_u32(&bar_full[s]), 1); mbar_init(smem_u32(&bar_empty[s]), nwarp); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  const int P = g.P;
  const int r_lo = rg.js - 3 + g.roff, r_hi = rg.je + 3 + g.roff;       // padded rows of the march
  const int nsteps = r_hi - r_lo + 1;
  // act_s[it]: the face row J = r-2 met at step it is part of the sweep (:868-869 domore_v, via `entry`); one
  // byte per step, filled once per CTA, so that the march reads it with a single LDS instead of a global load
  unsigned char *act_s = y_smem + sizeof(YStage) * YNS;
  {
    const int *__restrict__ ent = entry + k0 * g.R;
    const int rf_lo = rg.js - 1 + g.roff, rf_hi = rg.je + g.roff;       // face rows of the sweep
    for (int it = threadIdx.x; it < nsteps; it += blockDim.x) {
      const int rf = r_lo + it - 2;
      act_s[it] = (rf >= rf_lo && rf <= rf_hi && ent[rf]) ? 1 : 0;
    }
  }
  __syncthreads();
  const long long lev = g.plane * k0 + cbase;

  if (w == YW) {
    // ---- producer warp.  Everything it computes is warp-uniform (block indices, kernel parameters, the step
    // counter), and ONE elected lane issues the step's 6 + nact bulk copies back to back, so the addresses stay
    // in uniform registers: no per-lane serialisation loop (ELECT / R2UR / BRA.ANY per copy) in front of UBLKCP.
    const double *const msrc[6] = {hhv + lev, hcv + lev, gqv + lev, hlstv + lev, ihv + lev, mmv + cbase};
    const int mroff[6] = {2, 2, 2, 3, 3, 1};      // face r-2, cell r-3, mask product of cell r-1
    for (int it = 0; it < nsteps; it++) {
      const int s = it % YNS;
      mbar_wait_sleep(smem_u32(&bar_empty[s]), ((it / YNS) & 1) ^ 1);
      if (elect_one()) {
        const unsigned fb = smem_u32(&bar_full[s]);
        mbar_expect_tx(fb, (6 + nact) * YROWB);
        const unsigned sbase = smem_u32(&stg[s]);
        const int r = r_lo + it;
#pragma unroll
        for (int t = 0; t < 6; t++) bulk_g2s(sbase + t * YROWB, msrc[t] + (long long)P * max(r - mroff[t], 0), YROWB, fb);
        for (int t = 0; t < nact; t++)
          bulk_g2s(sbase + (6 + t) * YROWB, tt.dst[m0 + t] + lev + (long long)P * max(r, 0), YROWB, fb);
      }
      __syncwarp();
    }
    return;
  }
  if (w >= nwarp) return;

  // ---- consumer warp: tracers m0+2w and m0+2w+1 (n = 0, 1), column cbase + lane ----
  const int t0 = YTW * w;
  const bool two = (YTW > 1) && (t0 + 1 < nact);                          // the warp's second tracer exists
  const int ts[2] = {t0, two ? t0 + 1 : t0};                 // (a lone tracer is computed twice, stored once)
  const int i0 = cbase + lane - g.coff;
  const bool inr = (i0 >= rg.is && i0 <= rg.ie);
  // in place: the store of step it goes to row r_lo + it - 3; one running pointer, the second tracer at a fixed distance
  double *wp = tt.dst[m0 + ts[0]] + lev + lane + (long long)P * (r_lo - 3);
  const long long wd = tt.dst[m0 + ts[1]] - tt.dst[m0 + ts[0]];
  const double cu[2] = {tt.cu[m0 + ts[0]], tt.cu[m0 + ts[1]]};
  const bool any_cu = (cu[0] > 0.0) || (YTW > 1 && cu[1] > 0.0);
  unsigned bfull0 = smem_u32(&bar_full[0]), bempty0 = smem_u32(&bar_empty[0]);
  asm volatile("" : "+r"(bfull0), "+r"(bempty0));      // computed once (the conversion reads a special register)
  const unsigned sb = smem_u32(y_smem) + lane * 8;                      // this lane's column in stage 0, row 0
  const unsigned oT0 = (6 + ts[0]) * YROWB, dT = (ts[1] - ts[0]) * YROWB;
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
    mbar_wait(bfull0 + 8u * s, (it / YNS) & 1);
    const unsigned qa = sb + s * (unsigned)sizeof(YStage);
    const double Tn[2] = {lds_f64(qa + oT0), (YTW > 1) ? lds_f64(qa + oT0 + dT) : 0.0};
    const double hh2 = lds_f64(qa), hc2 = lds_f64(qa + YROWB), gq2 = lds_f64(qa + 2 * YROWB);
    const double hl = lds_f64(qa + 3 * YROWB), ih = lds_f64(qa + 4 * YROWB), mm1 = lds_f64(qa + 5 * YROWB);
    __syncwarp();
    if (lane == 0) mbar_arrive(bempty0 + 8u * s);      // the stage may be refilled
    const bool act2 = act_s[it] != 0;
    const bool out_row = (r - 3 >= rc_lo);
    double a0[2], A20[2], s1[2] = {0, 0};
#pragma unroll
    for (int n = 0; n < YTW; n++) {
      a0[n] = Tn[n] - T1[n];                               // face r-1|r
      A20[n] = 2.0 * fabs(a0[n]);
      if (SCHEME != SCHEME_PPM_H3) s1[n] = slope_f(a0[n], a1[n], A20[n], A21[n], Tn[n] - T2[n], mm1);    // cell r-1
    }
    // parabola of cell r-2: xL/xR = edge values, dA, a6  (PLM: xL = xR = T, dA = 0.5*slope, a6 unused)
    double xL[2], xR[2], dA[2], a6[2] = {0, 0};
    if (SCHEME == SCHEME_PLM) {
#pragma unroll
      for (int n = 0; n < YTW; n++) { xL[n] = T2[n]; xR[n] = T2[n]; dA[n] = 0.5 * s2[n]; }
    } else {
      double S[2], t1[2], dd[2], t2[2], e2[2] = {0, 0};
#pragma unroll
      for (int n = 0; n < YTW; n++) {
        if (SCHEME == SCHEME_PPM_CW) { e2[n] = edge_cw(T2[n], T1[n], s2[n], s1[n]); xL[n] = e3[n]; xR[n] = e2[n]; }
        else edges_h3(T3[n], T2[n], T1[n], xL[n], xR[n]);
        ppm_pcm(xL[n], xR[n], T2[n], pp2[n]);
        ppm_lim_a(xL[n], xR[n], T2[n], dA[n], S[n], t1[n], dd[n], t2[n]);
      }
      if (div6_tiny_pos(dd[0]) | (YTW > 1 && div6_tiny_pos(dd[YTW - 1]))) {   // denormal dA^2: rare
#pragma unroll
        for (int n = 0; n < YTW; n++) if (div6_tiny_pos(dd[n])) t2[n] = div6_int(dd[n]);
      }
      if (ppm_over(t1[0], t2[0]) | (YTW > 1 && ppm_over(t1[YTW - 1], t2[YTW - 1]))) {      // overshoot correction: rare
#pragma unroll
        for (int n = 0; n < YTW; n++) {
          double u = t1[n];
          asm volatile("" : "+d"(u));      // re-test here instead of carrying predicates through the hot path
          if (ppm_over(u, t2[n])) ppm_fix(xL[n], xR[n], T2[n], u, dA[n], S[n]);
        }
      }
#pragma unroll
      for (int n = 0; n < YTW; n++) { a6[n] = ppm_a6(T2[n], S[n]); e3[n] = e2[n]; }
    }
    // cell r-2 feeds its north face r-2 when hh2 >= 0 (kept for the next step) and its south face r-3 when hh3 < 0
    const bool south = (hh3 < 0.0);                        // same for both tracers
    const bool do_i = (__double_as_longlong(ih) != 0LL);   // Ihnew > 0 (it is never negative)
    double o[2];
#pragma unroll
    for (int n = 0; n < YTW; n++) {
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
      flp[n] = fl3;
      Fn3[n] = Fn2;
      pp2[n] = mm1 * a0[n] * a1[n];                        // cell r-1: mask*(Tp-Tc)*(Tc-Tm)
      T3[n] = T2[n]; T2[n] = T1[n]; T1[n] = Tn[n];
      a1[n] = a0[n]; A21[n] = A20[n]; s2[n] = s1[n];
    }
    if (any_cu) {      // warp-uniform and rare: a real branch (and one for both tracers), not predication
      if (cu[0] > 0.0) o[0] = underflow_rare(o[0], cu[0], inr);
      if (YTW > 1 && cu[1] > 0.0) o[YTW - 1] = underflow_rare(o[YTW - 1], cu[1], inr);
    }
    hh3 = hh2; hc3 = hc2; gq3 = gq2;
    act3 = act2;
    if (out_row) {
      wp[0] = o[0];
      if (YTW > 1 && two) wp[wd] = o[YTW - 1];
    }
    wp += P;
  }
}

void launch_y_fast(cudaStream_t s, int scheme, const Geom &g, Range rg, const TrTable &tt, const double *hh,
                   const double *hc, const double *gq, const double *hlst, const double *ihnew, const double *mm,
                   const int *entry, const int *domore_k) {
#ifdef YSWAP
  dim3 grid((tt.n + YT - 1) / YT, g.P / YCOLS, g.nk), block(32 * (YW + 1));
#else
  dim3 grid(g.P / YCOLS, (tt.n + YT - 1) / YT, g.nk), block(32 * (YW + 1));
#endif
  const size_t smem = sizeof(YStage) * YNS + ((g.R + 6 + 15) / 16) * 16;      // stages + one activity byte per step
#define YLAUNCH(S)                                                                                        \
  do {                                                                                                    \
    static size_t attr_smem = 0;      /* grows with the tallest plan seen (the activity bytes scale with the rows) */ \
    if (smem > attr_smem) { cudaFuncSetAttribute(k_y_ring<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); attr_smem = smem; } \
    k_y_ring<S><<<grid, block, smem, s>>>(g, rg, tt, hh, hc, gq, hlst, ihnew, mm, entry, domore_k);       \
  } while (0)
  if (scheme == SCHEME_PLM) YLAUNCH(SCHEME_PLM);
  else if (scheme == SCHEME_PPM_CW) YLAUNCH(SCHEME_PPM_CW);
  else YLAUNCH(SCHEME_PPM_H3);
#undef YLAUNCH
}
