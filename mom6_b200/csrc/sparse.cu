// Sparse tracer update for residual iterations (itt >= 2 of
// /root/reference/src/tracer/MOM_tracer_advect.F90:203-312).
//
// After the first trip uhr/vhr are non-zero only at the few faces where the
// flux limiter cut the transport (:495-500), so only the cells next to those
// faces change (do_i, :637 / :1030).  k_mass_update appends every such cell to a
// list; k_sparse_compute evaluates the reference update for (listed cell,
// tracer) pairs into the tracer's idle ping-pong buffer and k_sparse_scatter copies them
// back, so the sweep is in place and costs O(active cells) instead of a full pass.
// conc_underflow (:683-687 / :1062-1066) only needs the updated cells: every
// other cell of the range already went through it on the first trip.
#include "internal.cuh"
#include "kernels.h"

template <int SCHEME, int DIR>
__global__ void __launch_bounds__(128)
k_sparse_compute(Geom g, Range rg, TrTable tt, const int *__restrict__ list, const int *__restrict__ cnt_dev,
                 const double *__restrict__ hh, const double *__restrict__ cflv, const double *__restrict__ hlstv,
                 const double *__restrict__ ihv, const double *__restrict__ mmv, const int *__restrict__ entry,
                 DiagTable dg, const double *__restrict__ IareaT, double Idt, const int *__restrict__ tflags) {
  // The list length is read on the device: the host never waits for it (one grid of fixed size strides over the
  // list).  New values go to the tracer's IDLE ping-pong buffer at the cell's own index (tt.dst), so the staging
  // capacity is the whole field and the list cannot overflow: there is no dense fall-back to decide on the host.
  const int cnt = *cnt_dev;
  const int st = (DIR == 0) ? 1 : g.P;
  constexpr int W = (SCHEME == SCHEME_PPM_CW) ? 3 : 2;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < cnt; e += gridDim.x * blockDim.x) {
    const long long q = list[e];
    const int k0 = (int)(q / g.plane);
    const int rem = (int)(q - g.plane * k0);
    const int rj = rem / g.P;
    const long long q2 = rem;
    // faces of the cell: hi = q, lo = q - st.  In x every face of an entered row is swept; in y a face row may
    // be unflagged, then its flux is +0 (:1016-1018)
    bool act_hi = true, act_lo = true;
    if (DIR == 1) {
      const int j = rj - g.roff;
      act_hi = (j >= rg.js - 1 && j <= rg.je) && entry[k0 * g.R + rj];
      act_lo = (j - 1 >= rg.js - 1) && entry[k0 * g.R + rj - 1];
    }
    const double uh_hi = hh[q], cf_hi = cflv[q], uh_lo = hh[q - st], cf_lo = cflv[q - st];
    const double hl = hlstv[q], ih = ihv[q];
    double mm[5];
#pragma unroll
    for (int s = -2; s <= 2; s++) mm[s + 2] = mmv[q2 + s * st];
    // Flux diagnostics (:665-667,671-676 / :1051-1056,1070-1076).  Every listed cell has do_i true.  A face between two
    // listed cells is accumulated ONCE: by its low-side cell (as that cell's high face); a listed cell takes its low
    // face only when the cell below it is not updated (its tile was not visited by the mass update -- tflags bit 1
    // clear, its Ihnew is then stale -- or its Ihnew is zero).
    bool lo_mine = false;
    if (tflags) {
      const long long ql = q - st;
      const int k0l = k0, reml = rem - st;      // (same level: a listed cell is never in the first row / column of a plane)
      const int tl = (k0l * g.R + reml / g.P) * (g.P >> 5) + ((reml % g.P) >> 5);
      lo_mine = !((tflags[tl] & 2) && ihv[ql] > 0.0);
    }
    const double iarea = IareaT ? IareaT[q2] : 0.0;
    for (int m = blockIdx.y; m < tt.n; m += gridDim.y) {
      const double *__restrict__ tin = tt.src[m];
      double t[7];
#pragma unroll
      for (int s = -3; s <= 3; s++) t[s + 3] = (s >= -W && s <= W) ? tin[q + s * st] : 0.0;
      double f_hi = 0.0, f_lo = 0.0;
      if (act_hi) f_hi = face_flux<SCHEME>(&t[1], &mm[1], uh_hi, cf_hi);
      if (act_lo) f_lo = face_flux<SCHEME>(&t[0], &mm[0], uh_lo, cf_lo);
      double out = (t[3] * hl - (f_hi - f_lo)) * ih;
      if (tflags) {
        double *__restrict__ ad = dg.ad[m];
        double *__restrict__ axy = dg.axy[m];
        if (ad) {
          if (act_hi) ad[q] = ad[q] + f_hi * Idt;
          if (act_lo && lo_mine) ad[q - st] = ad[q - st] + f_lo * Idt;
        }
        if (axy) axy[q] = axy[q] - (f_hi - f_lo) * Idt * iarea;
      }
      const double cu = tt.cu[m];
      if (cu > 0.0 && fabs(out) < cu) out = 0.0;
      tt.dst[m][q] = out;
    }
  }
}

// tt.src = the idle buffers holding the staged values, tt.dst = the live buffers.  dirty (multi-GPU, nullable): per
// level and 32-column tile "a tracer cell changed since the last group pass" (the neighbours' pulls skip the rest).
__global__ void __launch_bounds__(128)
k_sparse_scatter(TrTable tt, const int *__restrict__ list, const int *__restrict__ cnt_dev,
                 unsigned char *__restrict__ dirty, long long plane, int P) {
  const int cnt = *cnt_dev;
  for (int e = blockIdx.x * blockDim.x + threadIdx.x; e < cnt; e += gridDim.x * blockDim.x) {
    const long long q = list[e];
    for (int m = blockIdx.y; m < tt.n; m += gridDim.y) tt.dst[m][q] = tt.src[m][q];
    if (dirty && blockIdx.y == 0) {
      const int k0 = (int)(q / plane), rem = (int)(q - plane * k0);
      const int rj = rem / P, ci = rem - rj * P;
      dirty[(plane >> 5) * k0 + (long long)rj * (P >> 5) + (ci >> 5)] = 1;
    }
  }
}

// tt: src = live buffers (read), dst = idle buffers (staging); the update is in place in the live buffers.
void launch_sparse_sweep(cudaStream_t s, int scheme, int dir, const Geom &g, Range rg, const TrTable &tt,
                         const int *list, const int *cnt_dev, int nsm, const double *hh, const double *cfl,
                         const double *hlst, const double *ihnew, const double *mm, const int *entry,
                         unsigned char *dirty, const DiagTable *dg, const double *IareaT, double Idt, const int *tflags) {
  dim3 grid(nsm * 2, tt.n < 8 ? tt.n : 8, 1);      // a residual trip lists 1e3-1e5 cells: enough threads, few idle CTAs
  static const DiagTable none = {};
  const DiagTable &dgt = dg ? *dg : none;
  const int *tf = dg ? tflags : nullptr;
#define SP(S, D) k_sparse_compute<S, D><<<grid, 128, 0, s>>>(g, rg, tt, list, cnt_dev, hh, cfl, hlst, ihnew, mm, entry, dgt, dg ? IareaT : nullptr, Idt, tf)
  if (scheme == SCHEME_PLM) { if (dir == 0) SP(SCHEME_PLM, 0); else SP(SCHEME_PLM, 1); }
  else if (scheme == SCHEME_PPM_CW) { if (dir == 0) SP(SCHEME_PPM_CW, 0); else SP(SCHEME_PPM_CW, 1); }
  else { if (dir == 0) SP(SCHEME_PPM_H3, 0); else SP(SCHEME_PPM_H3, 1); }
#undef SP
  TrTable back;
  back.n = tt.n;
  for (int m = 0; m < tt.n; m++) { back.src[m] = tt.dst[m]; back.dst[m] = const_cast<double *>(tt.src[m]); back.cu[m] = tt.cu[m]; }
  k_sparse_scatter<<<grid, 128, 0, s>>>(back, list, cnt_dev, dirty, g.plane, g.P);
}
