// Sparse tracer update for residual iterations (itt >= 2 of
// /root/reference/src/tracer/MOM_tracer_advect.F90:203-312).
//
// After the first trip uhr/vhr are non-zero only at the few faces where the
// flux limiter cut the transport (:495-500), so only the cells next to those
// faces change (do_i, :637 / :1030).  k_mass_update appends every such cell to a
// list; k_sparse_compute evaluates the reference update for (listed cell,
// tracer) pairs into a compact buffer and k_sparse_scatter writes them back, so
// the sweep is in place and costs O(active cells) instead of a full pass.
// conc_underflow (:683-687 / :1062-1066) only needs the updated cells: every
// other cell of the range already went through it on the first trip.
#include "internal.cuh"
#include "kernels.h"

template <int SCHEME, int DIR>
__global__ void __launch_bounds__(128)
k_sparse_compute(Geom g, Range rg, TrTable tt, const int *__restrict__ list, int cnt, int cap,
                 const double *__restrict__ hh, const double *__restrict__ cflv, const double *__restrict__ hlstv,
                 const double *__restrict__ ihv, const double *__restrict__ mmv, const int *__restrict__ entry,
                 double *__restrict__ newT) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= cnt) return;
  const int q32 = list[e];
  const long long q = q32;
  const int k0 = (int)(q / g.plane);
  const int rem = (int)(q - g.plane * k0);
  const int rj = rem / g.P;
  const long long q2 = rem;
  const int st = (DIR == 0) ? 1 : g.P;
  constexpr int W = (SCHEME == SCHEME_PPM_CW) ? 3 : 2;
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
  for (int m = blockIdx.y; m < tt.n; m += gridDim.y) {
    const double *__restrict__ tin = tt.src[m];
    double t[7];
#pragma unroll
    for (int s = -3; s <= 3; s++) t[s + 3] = (s >= -W && s <= W) ? tin[q + s * st] : 0.0;
    double f_hi = 0.0, f_lo = 0.0;
    if (act_hi) f_hi = face_flux<SCHEME>(&t[1], &mm[1], uh_hi, cf_hi);
    if (act_lo) f_lo = face_flux<SCHEME>(&t[0], &mm[0], uh_lo, cf_lo);
    double out = (t[3] * hl - (f_hi - f_lo)) * ih;
    const double cu = tt.cu[m];
    if (cu > 0.0 && fabs(out) < cu) out = 0.0;
    newT[(long long)m * cap + e] = out;
  }
}

__global__ void k_sparse_scatter(TrTable tt, const int *__restrict__ list, int cnt, int cap,
                                 const double *__restrict__ newT) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= cnt) return;
  const long long q = list[e];
  for (int m = blockIdx.y; m < tt.n; m += gridDim.y) tt.dst[m][q] = newT[(long long)m * cap + e];
}

void launch_sparse_sweep(cudaStream_t s, int scheme, int dir, const Geom &g, Range rg, const TrTable &tt,
                         const int *list, int cnt, int cap, const double *hh, const double *cfl, const double *hlst,
                         const double *ihnew, const double *mm, const int *entry, double *newT) {
  if (cnt <= 0) return;
  dim3 grid((cnt + 127) / 128, tt.n < 16 ? tt.n : 16, 1);
#define SP(S, D) k_sparse_compute<S, D><<<grid, 128, 0, s>>>(g, rg, tt, list, cnt, cap, hh, cfl, hlst, ihnew, mm, entry, newT)
  if (scheme == SCHEME_PLM) { if (dir == 0) SP(SCHEME_PLM, 0); else SP(SCHEME_PLM, 1); }
  else if (scheme == SCHEME_PPM_CW) { if (dir == 0) SP(SCHEME_PPM_CW, 0); else SP(SCHEME_PPM_CW, 1); }
  else { if (dir == 0) SP(SCHEME_PPM_H3, 0); else SP(SCHEME_PPM_H3, 1); }
#undef SP
  k_sparse_scatter<<<grid, 128, 0, s>>>(tt, list, cnt, cap, newT);
}
