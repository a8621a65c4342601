// Open-boundary branches of advect_x / advect_y
// (/root/reference/src/tracer/MOM_tracer_advect.F90:442-478,580-628 and :825-862,965-1014).
//
// With an associated OBC the sweeps run through these kernels instead of the streaming ones.  They follow the
// reference's own order of operations, one tracer at a time, on scratch fields:
//
//   Ttmp   = copy of the tracer (T_tmp, :436-440 / :820-822)
//   slope  = PLM slopes from the ORIGINAL tracer (:411-433 / :793-815)
//   per segment, in order: overwrite Ttmp on the exterior side of the segment's faces with the reservoir
//            value tres or OBC_inflow_conc, then recompute the three slopes around the face from Ttmp
//   hh2    = uhh with uhh := uhr on the segment faces that take a prescribed inflow (:590-626); the volume update
//            (uhr, hprev, hlst, Ihnew) uses hh2, the interior flux formula keeps the limited uhh
//   flux   = interior formula from Ttmp / slope, replaced by hh2 * (tres | OBC_inflow_conc) for the tracers of a
//            segment whose inflow condition held (cond bits written by k_obc_mass)
//
// Segment tables are tiny; every kernel walks them in the reference's order so that "the last assignment wins"
// exactly where the reference's loops overlap.  Regional configurations are small: this path trades speed for
// being a literal restatement.
#include "obc.cuh"

#define OTPB 128

__device__ __forceinline__ long long oidx(const Geom &g, int i, int j, int k0) {
  return (long long)(i + g.coff) + (long long)g.P * (j + g.roff) + g.plane * k0;
}
__device__ __forceinline__ double obc_tr_value(const ObcSeg &sg, const ObcTr &t, int i, int j, int k0) {
  if (!t.tres) return t.conc;
  // tres(IsdB:IedB, jsd:jed, nk) on E/W segments, (isd:ied, JsdB:JedB, nk) on N/S segments
  if (sg.is_ew) {
    const long long ni = sg.IedB - sg.IsdB + 1, nj = sg.jed - sg.jsd + 1;
    return t.tres[(i - sg.IsdB) + ni * ((j - sg.jsd) + nj * (long long)k0)];
  }
  const long long ni = sg.ied - sg.isd + 1, nj = sg.JedB - sg.JsdB + 1;
  return t.tres[(i - sg.isd) + ni * ((j - sg.JsdB) + nj * (long long)k0)];
}
// index of face (i,j,k0) in the segment's cond array
__device__ __forceinline__ long long obc_cond_idx(const ObcSeg &sg, int i, int j, int k0) {
  if (sg.is_ew) { const long long nj = sg.jed - sg.jsd + 1; return sg.cond_off + (j - sg.jsd) + nj * (long long)k0; }
  const long long ni = sg.ied - sg.isd + 1, nj = sg.JedB - sg.JsdB + 1;
  return sg.cond_off + (i - sg.isd) + ni * ((j - sg.JsdB) + nj * (long long)k0);
}

// PLM slopes of one tracer from its original values (:411-433 / :793-815)
template <int DIR>
__global__ void k_obc_slope(Geom g, Range rg, int sten, const double *__restrict__ T, const double *__restrict__ mm,
                            double *__restrict__ slope, const int *__restrict__ domore_k) {
  const int ci = blockIdx.x * blockDim.x + threadIdx.x, rj = blockIdx.y, k0 = blockIdx.z;
  if (ci >= g.P || domore_k[k0] <= 0) return;
  const int i = ci - g.coff, j = rj - g.roff;
  const bool in = (DIR == 0) ? (i >= rg.is - sten && i <= rg.ie + sten && j >= rg.js && j <= rg.je)
                             : (i >= rg.is && i <= rg.ie && j >= rg.js - sten && j <= rg.je + sten);
  if (!in) return;
  const int st = (DIR == 0) ? 1 : g.P;
  const long long q = (long long)ci + (long long)g.P * rj + g.plane * k0, q2 = (long long)ci + (long long)g.P * rj;
  slope[q] = plm_slope(T[q + st], T[q], T[q - st], mm[q2]);
}

// T_tmp overrides of one segment for tracer m, then the three slopes around its faces (:442-478 / :825-862)
template <int DIR>
__global__ void k_obc_override(Geom g, Range rg, ObcSeg sg, const ObcTr *__restrict__ trs, int m,
                               double *__restrict__ Ttmp, double *__restrict__ slope, const double *__restrict__ mm,
                               const int *__restrict__ entry, const int *__restrict__ domore_k) {
  const int a = blockIdx.x * blockDim.x + threadIdx.x, k0 = blockIdx.y;
  if (domore_k[k0] <= 0) return;
  int i, j;
  if (DIR == 0) {
    j = sg.jsd + a; i = sg.IsdB;
    if (j > sg.jed || j < rg.js || j > rg.je) return;
    if (!entry[k0 * g.R + j + g.roff]) return;                 // rows without domore_u are skipped (:407)
  } else {
    i = sg.isd + a; j = sg.JsdB;
    if (i > sg.ied || i < rg.is || i > rg.ie) return;
  }
  const int st = (DIR == 0) ? 1 : g.P;
  const long long q = oidx(g, i, j, k0), q2 = oidx(g, i, j, 0);
  const bool low = (DIR == 0) ? (sg.dir == OBC_DIR_W) : (sg.dir == OBC_DIR_S);   // exterior cell = the face's own index
  for (int e = 0; e < sg.ntseg; e++) {
    const ObcTr t = trs[sg.tr_off + e];
    if (t.m != m) continue;
    Ttmp[low ? q : q + st] = obc_tr_value(sg, t, i, j, k0);
  }
  for (int d = -1; d <= 1; d++) {
    const long long c = q + d * st;
    slope[c] = plm_slope(Ttmp[c + st], Ttmp[c], Ttmp[c - st], mm[q2 + d * st]);
  }
}

// uhh := uhr on the segment faces that take a prescribed inflow (:580-628 / :965-1014); blk 1: directional test,
// blk 2: inflow from a masked cell on an open (not specified) segment.  Writes the cond bits the flux kernel reads.
template <int DIR>
__global__ void k_obc_mass(Geom g, Range rg, ObcSeg sg, int blk, const double *__restrict__ res,
                           double *__restrict__ hh2, const double *__restrict__ maskT, unsigned char *__restrict__ cond,
                           const int *__restrict__ entry, const int *__restrict__ domore_k) {
  const int a = blockIdx.x * blockDim.x + threadIdx.x, k0 = blockIdx.y;
  if (domore_k[k0] <= 0) return;
  int i, j;
  if (DIR == 0) {
    j = sg.jsd + a; i = sg.IsdB;
    if (j > sg.jed || j < rg.js || j > rg.je || i < rg.is - 1 || i > rg.ie) return;
  } else {
    const int ni = sg.ied - sg.isd + 1;
    i = sg.isd + a % ni; j = sg.JsdB + a / ni;
    if (j > sg.JedB || j < rg.js - 1 || j > rg.je || i < rg.is || i > rg.ie) return;
  }
  if (!entry[k0 * g.R + j + g.roff]) return;                   // inside "if (domore_u(j,k))" / "if (domore_v(J,k))"
  const int st = (DIR == 0) ? 1 : g.P;
  const long long q = oidx(g, i, j, k0), q2 = oidx(g, i, j, 0);
  const double r = res[q];
  bool c;
  if (blk == 1) {
    c = (DIR == 0) ? ((r > 0.0 && sg.dir == OBC_DIR_W) || (r < 0.0 && sg.dir == OBC_DIR_E))
                   : ((r > 0.0 && sg.dir == OBC_DIR_S) || (r < 0.0 && sg.dir == OBC_DIR_N));
  } else {
    c = (r > 0.0 && maskT[q2] < 0.5) || (r < 0.0 && maskT[q2 + st] < 0.5);
  }
  if (!c) return;
  hh2[q] = r;
  cond[obc_cond_idx(sg, i, j, k0)] |= (unsigned char)blk;
}

// interior flux through the face between cells a and a+st from T_tmp and the slope field (:517-577 / :902-963)
template <int SCHEME>
__device__ __forceinline__ double obc_flux_interior(const double *__restrict__ T, const double *__restrict__ sl,
                                                    const double *__restrict__ mm, long long qa, long long q2a, int st,
                                                    double uhh, double cfl) {
  const long long u = (uhh >= 0.0) ? qa : qa + st, u2 = (uhh >= 0.0) ? q2a : q2a + st;
  if (SCHEME == SCHEME_PLM) {
    if (uhh >= 0.0) return uhh * (T[u] + 0.5 * sl[u] * (1. - cfl));
    return uhh * (T[u] - 0.5 * sl[u] * (1. - cfl));
  }
  const double Tp = T[u + st], Tc = T[u], Tm = T[u - st];
  double aL, aR;
  if (SCHEME == SCHEME_PPM_H3) {
    aL = div6(5. * Tc + (2. * Tm - Tp));
    aL = gmax2(gmin2(Tc, Tm), aL); aL = gmin2(gmax2(Tc, Tm), aL);
    aR = div6(5. * Tc + (2. * Tp - Tm));
    aR = gmax2(gmin2(Tc, Tp), aR); aR = gmin2(gmax2(Tc, Tp), aR);
  } else {
    aL = 0.5 * ((Tm + Tc) + div3(sl[u - st] - sl[u]));
    aR = 0.5 * ((Tc + Tp) + div3(sl[u] - sl[u + st]));
  }
  const double dA = aR - aL, mA = 0.5 * (aR + aL);
  const double t1 = dA * (Tc - mA), t2 = div6(dA * dA);
  if (mm[u2] * (Tp - Tc) * (Tc - Tm) <= 0.) { aL = Tc; aR = Tc; }
  else if (t1 > t2) { aL = 3. * Tc - 2. * aR; }
  else if (t1 < -t2) { aR = 3. * Tc - 2. * aL; }
  const double a6 = 6. * Tc - 3. * (aR + aL);
  const double c23 = 2. / 3.;
  if (uhh >= 0.0) return uhh * (aR - 0.5 * cfl * ((aR - aL) - a6 * (1. - c23 * cfl)));
  return uhh * (aL + 0.5 * cfl * ((aR - aL) + a6 * (1. - c23 * cfl)));
}

// prescribed flux of tracer m through face (i,j,k0), if a segment's inflow condition held there; the segment
// loops run in the reference's order (block 1 over all segments, then block 2), the last assignment wins
template <int DIR>
__device__ __forceinline__ bool obc_flux_override(const ObcSeg *__restrict__ segs, int nseg, const ObcTr *__restrict__ trs,
                                                  const unsigned char *__restrict__ cond, int m, int i, int j, int k0,
                                                  double *val) {
  bool found = false;
  for (int blk = 1; blk <= 2; blk++) for (int n = 0; n < nseg; n++) {
    const ObcSeg sg = segs[n];
    if (!sg.has_reg || ((DIR == 0) ? !sg.is_ew : !sg.is_ns)) continue;
    const bool on = (DIR == 0) ? (i == sg.IsdB && j >= sg.jsd && j <= sg.jed)
                               : (j >= sg.JsdB && j <= sg.JedB && i >= sg.isd && i <= sg.ied);
    if (!on) continue;
    if (!(cond[obc_cond_idx(sg, i, j, k0)] & blk)) continue;
    for (int e = 0; e < sg.ntseg; e++) {
      const ObcTr t = trs[sg.tr_off + e];
      if (t.m == m) { *val = obc_tr_value(sg, t, i, j, k0); found = true; }
    }
  }
  return found;
}

// fluxes of both faces of a cell, tracer update, conc_underflow and the 3-D diagnostics of ONE tracer
// (:517-577,652-687 / :902-963,1043-1066); fx receives the flux through the cell's east / north face (for ad2d)
template <int SCHEME, int DIR>
__global__ void __launch_bounds__(OTPB)
k_obc_flux_update(Geom g, Range rg, TracerDev tr, int m, int src, const double *__restrict__ Ttmp,
                  const double *__restrict__ slope, const double *__restrict__ hh, const double *__restrict__ hh2,
                  const double *__restrict__ cflv, const double *__restrict__ hlstv, const double *__restrict__ ihnewv,
                  const double *__restrict__ mmv, const double *__restrict__ IareaT, const int *__restrict__ entry,
                  const int *__restrict__ domore_k, double Idt, const ObcSeg *__restrict__ segs, int nseg,
                  const ObcTr *__restrict__ trs, const unsigned char *__restrict__ cond, double *__restrict__ fx) {
  const int ci = blockIdx.x * blockDim.x + threadIdx.x, rj = blockIdx.y, k0 = blockIdx.z;
  if (ci >= g.P) return;
  const int i = ci - g.coff, j = rj - g.roff;
  if (i < g.isd || i > g.ied || j < g.jsd || j > g.jed) return;
  const long long q = (long long)ci + (long long)g.P * rj + g.plane * k0, q2 = (long long)ci + (long long)g.P * rj;
  const int st = (DIR == 0) ? 1 : g.P;
  const bool lev = domore_k[k0] > 0;
  const bool in_rng = (i >= rg.is && i <= rg.ie && j >= rg.js && j <= rg.je);
  const bool hi_ok = (DIR == 0) ? (i >= rg.is - 1 && i <= rg.ie && j >= rg.js && j <= rg.je)
                                : (i >= rg.is && i <= rg.ie && j >= rg.js - 1 && j <= rg.je);
  const bool act_hi = lev && hi_ok && entry[k0 * g.R + rj];
  const bool act_lo = (DIR == 0) ? act_hi : (lev && in_rng && entry[k0 * g.R + rj - 1]);
  const double ih = (lev && in_rng) ? ihnewv[q] : 0.0;
  const bool do_i = ih > 0.0;
  const bool nxt_in = (DIR == 0) ? (i + 1 >= rg.is && i + 1 <= rg.ie && j >= rg.js && j <= rg.je)
                                 : (i >= rg.is && i <= rg.ie && j + 1 >= rg.js && j + 1 <= rg.je);
  const bool do_nxt = lev && nxt_in && (ihnewv[q + st] > 0.0);
  const bool diag_hi = act_hi && (do_i || do_nxt);
  const double *__restrict__ tin = tr.buf[src];
  double *__restrict__ tout = tr.buf[1 - src];
  const double Tc = tin[q];
  double out = Tc, f_hi = 0.0;
  if (do_i || diag_hi) {
    double f_lo = 0.0, v;
    if (act_hi) {
      f_hi = obc_flux_interior<SCHEME>(Ttmp, slope, mmv, q, q2, st, hh[q], cflv[q]);
      if (obc_flux_override<DIR>(segs, nseg, trs, cond, m, i, j, k0, &v)) f_hi = hh2[q] * v;
    }
    if (do_i) {
      if (act_lo) {
        f_lo = obc_flux_interior<SCHEME>(Ttmp, slope, mmv, q - st, q2 - st, st, hh[q - st], cflv[q - st]);
        if (obc_flux_override<DIR>(segs, nseg, trs, cond, m, (DIR == 0) ? i - 1 : i, (DIR == 0) ? j : j - 1, k0, &v))
          f_lo = hh2[q - st] * v;
      }
      out = (Tc * hlstv[q] - (f_hi - f_lo)) * ih;
      if (tr.adv_xy) tr.adv_xy[q] = tr.adv_xy[q] - (f_hi - f_lo) * Idt * IareaT[q2];
    }
    double *ad = (DIR == 0) ? tr.ad_x : tr.ad_y;
    if (ad && diag_hi) ad[q] = ad[q] + f_hi * Idt;
  }
  if (fx) fx[q] = diag_hi ? f_hi : 0.0;
  const double cu = tr.conc_underflow;
  if (lev && in_rng && cu > 0.0 && fabs(out) < cu) out = 0.0;
  tout[q] = out;
}

// ad2d_x / ad2d_y of one tracer: the k-sum of the face fluxes kept in fx, in level order (:691-699 / :1077-1084)
template <int DIR>
__global__ void k_obc_ad2d(Geom g, Range rg, double *__restrict__ a2, const double *__restrict__ fx,
                           const double *__restrict__ ihnewv, const int *__restrict__ entry,
                           const int *__restrict__ domore_k, double Idt) {
  const int ci = blockIdx.x * blockDim.x + threadIdx.x, rj = blockIdx.y;
  if (ci >= g.P) return;
  const int i = ci - g.coff, j = rj - g.roff;
  const bool hi_ok = (DIR == 0) ? (i >= rg.is - 1 && i <= rg.ie && j >= rg.js && j <= rg.je)
                                : (i >= rg.is && i <= rg.ie && j >= rg.js - 1 && j <= rg.je);
  if (!hi_ok) return;
  const int st = (DIR == 0) ? 1 : g.P;
  const bool cur_in = (i >= rg.is && i <= rg.ie && j >= rg.js && j <= rg.je);
  const bool nxt_in = (DIR == 0) ? (i + 1 <= rg.ie) : (j + 1 <= rg.je);
  const long long q2 = (long long)ci + (long long)g.P * rj;
  double acc = a2[q2];
  for (int k0 = 0; k0 < g.nk; k0++) {
    if (domore_k[k0] <= 0 || !entry[k0 * g.R + rj]) continue;
    const long long q = q2 + g.plane * k0;
    const bool d = (cur_in && ihnewv[q] > 0.0) || (nxt_in && ihnewv[q + st] > 0.0);
    if (!d) continue;
    acc = acc + fx[q] * Idt;
  }
  a2[q2] = acc;
}

// ---- launchers -------------------------------------------------------------------------------------------
void launch_obc_slope(cudaStream_t s, int dir, const Geom &g, Range rg, int sten, const double *T, const double *mm,
                      double *slope, const int *domore_k) {
  dim3 grid((g.P + OTPB - 1) / OTPB, g.R, g.nk);
  if (dir == 0) k_obc_slope<0><<<grid, OTPB, 0, s>>>(g, rg, sten, T, mm, slope, domore_k);
  else k_obc_slope<1><<<grid, OTPB, 0, s>>>(g, rg, sten, T, mm, slope, domore_k);
}
static int seg_len(const ObcSeg &sg, int dir, bool faces) {
  if (dir == 0) return sg.jed - sg.jsd + 1;
  return (sg.ied - sg.isd + 1) * (faces ? (sg.JedB - sg.JsdB + 1) : 1);
}
void launch_obc_override(cudaStream_t s, int dir, const Geom &g, Range rg, const ObcSeg &sg, const ObcTr *trs, int m,
                         double *Ttmp, double *slope, const double *mm, const int *entry, const int *domore_k) {
  const int n = seg_len(sg, dir, false);
  if (n <= 0) return;
  dim3 grid((n + OTPB - 1) / OTPB, g.nk);
  if (dir == 0) k_obc_override<0><<<grid, OTPB, 0, s>>>(g, rg, sg, trs, m, Ttmp, slope, mm, entry, domore_k);
  else k_obc_override<1><<<grid, OTPB, 0, s>>>(g, rg, sg, trs, m, Ttmp, slope, mm, entry, domore_k);
}
void launch_obc_mass(cudaStream_t s, int dir, const Geom &g, Range rg, const ObcSeg &sg, int blk, const double *res,
                     double *hh2, const double *maskT, unsigned char *cond, const int *entry, const int *domore_k) {
  const int n = seg_len(sg, dir, true);
  if (n <= 0) return;
  dim3 grid((n + OTPB - 1) / OTPB, g.nk);
  if (dir == 0) k_obc_mass<0><<<grid, OTPB, 0, s>>>(g, rg, sg, blk, res, hh2, maskT, cond, entry, domore_k);
  else k_obc_mass<1><<<grid, OTPB, 0, s>>>(g, rg, sg, blk, res, hh2, maskT, cond, entry, domore_k);
}
void launch_obc_flux_update(cudaStream_t s, int scheme, int dir, const Geom &g, Range rg, const TracerDev &tr, int m,
                            int src, const double *Ttmp, const double *slope, const double *hh, const double *hh2,
                            const double *cfl, const double *hlst, const double *ihnew, const double *mm,
                            const double *IareaT, const int *entry, const int *domore_k, double Idt,
                            const ObcSeg *segs, int nseg, const ObcTr *trs, const unsigned char *cond, double *fx) {
  dim3 grid((g.P + OTPB - 1) / OTPB, g.R, g.nk);
#define OFU(S, D) k_obc_flux_update<S, D><<<grid, OTPB, 0, s>>>(g, rg, tr, m, src, Ttmp, slope, hh, hh2, cfl, hlst, ihnew, \
                                                                 mm, IareaT, entry, domore_k, Idt, segs, nseg, trs, cond, fx)
  if (scheme == SCHEME_PLM) { if (dir == 0) OFU(SCHEME_PLM, 0); else OFU(SCHEME_PLM, 1); }
  else if (scheme == SCHEME_PPM_CW) { if (dir == 0) OFU(SCHEME_PPM_CW, 0); else OFU(SCHEME_PPM_CW, 1); }
  else { if (dir == 0) OFU(SCHEME_PPM_H3, 0); else OFU(SCHEME_PPM_H3, 1); }
#undef OFU
}
void launch_obc_ad2d(cudaStream_t s, int dir, const Geom &g, Range rg, double *a2, const double *fx,
                     const double *ihnew, const int *entry, const int *domore_k, double Idt) {
  dim3 grid((g.P + OTPB - 1) / OTPB, g.R, 1);
  if (dir == 0) k_obc_ad2d<0><<<grid, OTPB, 0, s>>>(g, rg, a2, fx, ihnew, entry, domore_k, Idt);
  else k_obc_ad2d<1><<<grid, OTPB, 0, s>>>(g, rg, a2, fx, ihnew, entry, domore_k, Idt);
}
