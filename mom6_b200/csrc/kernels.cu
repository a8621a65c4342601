// Hand-written sm_100a kernels for MOM6 advect_tracer (unfused reference-order
// path).  Arithmetic follows /root/reference/src/tracer/MOM_tracer_advect.F90
// operation for operation; build with -fmad=false (no FMA contraction).
//
// Kernel inventory:
//   k_init_residuals   K0  :152-178   uhr/vhr/hprev from h_end/uhtr/vhtr
//   k_neglect          K1  :182-188   uh_neglect / vh_neglect (+ mask products)
//   k_rescan_flags     K3  :215-225   domore_u/domore_v row scans (warp ballots)
//   k_domore_k         K7  :229-231,265-267,287-289
//   k_flags_begin          :394,407-408 / :789,868-869  entry flags of a sweep
//   k_mass_flux<DIR>       :485-514 / :871-900  uhh/vhh, CFL, limiter flags
//   k_mass_update<DIR>     :632-649 / :1021-1040  uhr/vhr, hprev, hlst, Ihnew
//   k_tracer_sweep<SCHEME,DIR>  K4/K5/K6 :517-577,652-687 / :902-963,1043-1066
//   k_ad2d<SCHEME,DIR>     :691-699 / :1078-1084  k-ordered 2-D flux sums
//   k_halo_apply       C1  :206  list-driven halo fill (cyclic / fold / closed)
//   k_repack_in/out        Fortran layout <-> padded HBM layout
#include "internal.cuh"
#include "kernels.h"

#define TPB 128

__device__ __forceinline__ long long gidx(const Geom &g, int ci, int rj, int k0) {
  return (long long)ci + (long long)g.P * rj + g.plane * k0;
}

// ---------------------------------------------------------------------------
// Fortran layout <-> padded layout.  src: (ilo:ied, jlo:jed, nk) contiguous.
// ---------------------------------------------------------------------------
__global__ void k_repack_in(Geom g, double *__restrict__ dst, const double *__restrict__ src,
                            int ilo, int jlo, int nlev) {
  const int ni = g.ied - ilo + 1, nj = g.jed - jlo + 1;
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  int r = blockIdx.y, k0 = blockIdx.z;
  if (c >= ni || r >= nj || k0 >= nlev) return;
  dst[gidx(g, c + ilo + g.coff, r + jlo + g.roff, k0)] = src[(long long)c + (long long)ni * (r + (long long)nj * k0)];
}
__global__ void k_repack_out(Geom g, double *__restrict__ dst, const double *__restrict__ src,
                             int ilo, int jlo, int nlev) {
  const int ni = g.ied - ilo + 1, nj = g.jed - jlo + 1;
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  int r = blockIdx.y, k0 = blockIdx.z;
  if (c >= ni || r >= nj || k0 >= nlev) return;
  dst[(long long)c + (long long)ni * (r + (long long)nj * k0)] = src[gidx(g, c + ilo + g.coff, r + jlo + g.roff, k0)];
}

void launch_repack_in(cudaStream_t s, const Geom &g, double *dst, const double *src, int ilo, int jlo, int nlev) {
  const int ni = g.ied - ilo + 1, nj = g.jed - jlo + 1;
  dim3 grid((ni + TPB - 1) / TPB, nj, nlev);
  k_repack_in<<<grid, TPB, 0, s>>>(g, dst, src, ilo, jlo, nlev);
}
void launch_repack_out(cudaStream_t s, const Geom &g, double *dst, const double *src, int ilo, int jlo, int nlev) {
  const int ni = g.ied - ilo + 1, nj = g.jed - jlo + 1;
  dim3 grid((ni + TPB - 1) / TPB, nj, nlev);
  k_repack_out<<<grid, TPB, 0, s>>>(g, dst, src, ilo, jlo, nlev);
}

// ---------------------------------------------------------------------------
// K1: uh_neglect, vh_neglect (:182-188) and the per-cell mask products
//     mmx(i,j) = mask2dCu(I,j)*mask2dCu(I-1,j), mmy(i,j) = mask2dCv(i,J)*mask2dCv(i,J-1)
// ---------------------------------------------------------------------------
__global__ void k_neglect(Geom g, const double *__restrict__ areaT, const double *__restrict__ mCu,
                          const double *__restrict__ mCv, double H_subroundoff,
                          double *__restrict__ uh_neglect, double *__restrict__ vh_neglect,
                          double *__restrict__ mmx, double *__restrict__ mmy) {
  int ci = blockIdx.x * blockDim.x + threadIdx.x, rj = blockIdx.y;
  if (ci >= g.P) return;
  int i = ci - g.coff, j = rj - g.roff;
  long long q = gidx(g, ci, rj, 0);
  if (i >= g.isd && i <= g.ied && j >= g.jsd && j <= g.jed) {
    if (i <= g.ied - 1) uh_neglect[q] = H_subroundoff * gmin2(areaT[q], areaT[q + 1]);
    if (j <= g.jed - 1) vh_neglect[q] = H_subroundoff * gmin2(areaT[q], areaT[q + g.P]);
    if (i - 1 >= g.IsdB) mmx[q] = mCu[q] * mCu[q - 1];
    if (j - 1 >= g.JsdB) mmy[q] = mCv[q] * mCv[q - g.P];
  }
}
void launch_neglect(cudaStream_t s, const Geom &g, const double *areaT, const double *mCu, const double *mCv,
                    double Hsub, double *uhn, double *vhn, double *mmx, double *mmy) {
  dim3 grid((g.P + TPB - 1) / TPB, g.R, 1);
  k_neglect<<<grid, TPB, 0, s>>>(g, areaT, mCu, mCv, Hsub, uhn, vhn, mmx, mmy);
}

// ---------------------------------------------------------------------------
// K0: prologue :152-178.  Whole padded arrays are written (halos zeroed).
// ---------------------------------------------------------------------------
__global__ void k_init_residuals(Geom g, const double *__restrict__ h_end, const double *__restrict__ uhtr,
                                 const double *__restrict__ vhtr, const double *__restrict__ vol_prev,
                                 const double *__restrict__ areaT, double *__restrict__ uhr,
                                 double *__restrict__ vhr, double *__restrict__ hprev) {
  int ci = blockIdx.x * blockDim.x + threadIdx.x, rj = blockIdx.y, k0 = blockIdx.z;
  if (ci >= g.P) return;
  int i = ci - g.coff, j = rj - g.roff;
  long long q = gidx(g, ci, rj, k0);
  const bool in_i = (i >= g.isc && i <= g.iec), in_j = (j >= g.jsc && j <= g.jec);
  // uhr(I,j) = uhtr for I=is-1..ie, j=js..je ; vhr(i,J) = vhtr for i=is..ie, J=js-1..je
  double u = 0.0, v = 0.0, h = 0.0;
  if (in_j && i >= g.isc - 1 && i <= g.iec) u = uhtr[q];
  if (in_i && j >= g.jsc - 1 && j <= g.jec) v = vhtr[q];
  if (in_i && in_j) {
    if (vol_prev == nullptr) {
      double um = uhtr[q - 1], vm = vhtr[q - g.P];
      double ah = areaT[gidx(g, ci, rj, 0)] * h_end[q];
      h = gmax2(0.0, ah + ((u - um) + (v - vm)));
      h = h + gmax2(0.0, 1.0e-13 * h - ah);
    } else {
      h = vol_prev[q];
    }
  }
  uhr[q] = u; vhr[q] = v; hprev[q] = h;
}
void launch_init_residuals(cudaStream_t s, const Geom &g, const double *h_end, const double *uhtr,
                           const double *vhtr, const double *vol_prev, const double *areaT,
                           double *uhr, double *vhr, double *hprev) {
  dim3 grid((g.P + TPB - 1) / TPB, g.R, g.nk);
  k_init_residuals<<<grid, TPB, 0, s>>>(g, h_end, uhtr, vhtr, vol_prev, areaT, uhr, vhr, hprev);
}

// ---------------------------------------------------------------------------
// K3: domore_u / domore_v rescan, :215-225.  One CTA per (row, level); the row
// is scanned with coalesced loads and reduced with __syncthreads_or.
// Flags are int arrays indexed [k0*R + rj].
// ---------------------------------------------------------------------------
__global__ void k_rescan_flags(Geom g, const double *__restrict__ uhr, const double *__restrict__ vhr,
                               const int *__restrict__ domore_k, int *__restrict__ domore_u,
                               int *__restrict__ domore_v, int isv, int iev, int jsv, int jev, int stencil) {
  const int rj = blockIdx.x, k0 = blockIdx.y;
  if (domore_k[k0] <= 0) return;
  const int j = rj - g.roff;
  const long long base = (long long)g.P * rj + g.plane * k0;
  const bool scan_u = (j >= jsv && j <= jev) && !domore_u[k0 * g.R + rj];
  const bool scan_v = (j >= jsv + stencil - 1 && j <= jev - stencil) && !domore_v[k0 * g.R + rj];
  int fu = 0, fv = 0;
  if (scan_u) {
    for (int i = isv + stencil - 1 + threadIdx.x; i <= iev - stencil; i += blockDim.x)
      if (uhr[base + i + g.coff] != 0.0) fu = 1;
  }
  if (scan_v) {
    for (int i = isv + stencil + threadIdx.x; i <= iev - stencil; i += blockDim.x)
      if (vhr[base + i + g.coff] != 0.0) fv = 1;
  }
  fu = __syncthreads_or(fu);
  fv = __syncthreads_or(fv);
  if (threadIdx.x == 0) {
    if (scan_u && fu) domore_u[k0 * g.R + rj] = 1;
    if (scan_v && fv) domore_v[k0 * g.R + rj] = 1;
  }
}
void launch_rescan(cudaStream_t s, const Geom &g, const double *uhr, const double *vhr, const int *domore_k,
                   int *domore_u, int *domore_v, int isv, int iev, int jsv, int jev, int stencil) {
  dim3 grid(g.R, g.nk, 1);
  k_rescan_flags<<<grid, TPB, 0, s>>>(g, uhr, vhr, domore_k, domore_u, domore_v, isv, iev, jsv, jev, stencil);
}

// K7: domore_k(k) = any(domore_u(ju0:ju1,k)) .or. any(domore_v(jv0:jv1,k)), only for
// levels with domore_k(k) > 0 on entry (:229-231, :265-267, :287-289).
__global__ void k_domore_k(Geom g, int *__restrict__ domore_k, const int *__restrict__ domore_u,
                           const int *__restrict__ domore_v, int ju0, int ju1, int jv0, int jv1) {
  const int k0 = blockIdx.x;
  if (domore_k[k0] <= 0) return;   // uniform per CTA
  int f = 0;
  for (int j = ju0 + threadIdx.x; j <= ju1; j += blockDim.x) f |= domore_u[k0 * g.R + j + g.roff];
  for (int j = jv0 + threadIdx.x; j <= jv1; j += blockDim.x) f |= domore_v[k0 * g.R + j + g.roff];
  f = __syncthreads_or(f);
  if (threadIdx.x == 0) domore_k[k0] = f ? 1 : 0;
}
void launch_domore_k(cudaStream_t s, const Geom &g, int *domore_k, const int *domore_u, const int *domore_v,
                     int ju0, int ju1, int jv0, int jv1) {
  k_domore_k<<<g.nk, TPB, 0, s>>>(g, domore_k, domore_u, domore_v, ju0, ju1, jv0, jv1);
}

// Sum of domore_k over levels -> *out (the local part of the :303-307 test).
__global__ void k_sum_domore_k(const int *__restrict__ domore_k, int nk, int *__restrict__ out) {
  int s = 0;
  for (int k = threadIdx.x; k < nk; k += blockDim.x) s += domore_k[k];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
  __shared__ int ws[32];
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) { int t = 0; for (int w = 0; w < (blockDim.x + 31) / 32; w++) t += ws[w]; *out = t; }
}
void launch_sum_domore_k(cudaStream_t s, const int *domore_k, int nk, int *out) {
  k_sum_domore_k<<<1, TPB, 0, s>>>(domore_k, nk, out);
}

// Entry flags of a sweep: entry(r,k) = domore_k(k)>0 .and. lo<=r<=hi .and. domore(r,k);
// rows that will be processed have their flag cleared (:407-408 / :868-869).
__global__ void k_flags_begin(Geom g, const int *__restrict__ domore_k, int *__restrict__ domore,
                              int *__restrict__ entry, int lo, int hi) {
  int rj = blockIdx.x * blockDim.x + threadIdx.x, k0 = blockIdx.y;
  if (rj >= g.R) return;
  int r = rj - g.roff;
  int e = (domore_k[k0] > 0) && (r >= lo) && (r <= hi) && (domore[k0 * g.R + rj] != 0);
  entry[k0 * g.R + rj] = e;
  if (e) domore[k0 * g.R + rj] = 0;
}
void launch_flags_begin(cudaStream_t s, const Geom &g, const int *domore_k, int *domore, int *entry, int lo, int hi) {
  dim3 grid((g.R + TPB - 1) / TPB, g.nk, 1);
  k_flags_begin<<<grid, TPB, 0, s>>>(g, domore_k, domore, entry, lo, hi);
}

// ---------------------------------------------------------------------------
// Mass fluxes of one sweep: uhh/vhh, CFL and the limiter flag
// (:485-514 for x, :871-900 for y).  DIR 0 = x (face stride 1), 1 = y (stride P).
// Written for every point of the padded array (0 where the face is not part of
// the sweep) so that the kernels that follow need no flag logic of their own.
// ---------------------------------------------------------------------------
// Limited transport through one face and its CFL number (:485-514 / :871-900); ur = residual transport of the
// face if its row is swept, else 0.  Writes hh, cfl and the per-face factors of the flux formulas.
template <int DIR>
__device__ __forceinline__ void mass_flux_face(const Geom &g, long long q, long long q2, int fl, bool swept, double ur,
                                               const double *__restrict__ res, const double *__restrict__ hprev,
                                               const double *__restrict__ areaT, double min_h, int *__restrict__ domore,
                                               double *__restrict__ hh, double *__restrict__ cfl,
                                               double *__restrict__ hcv, double *__restrict__ gv, int plm) {
  const int st = (DIR == 0) ? 1 : g.P;
  double uh = 0.0, cf = 0.0;
  if (swept) {
    const double hm = hprev[q], hp = hprev[q + st];   // cells on the low / high side of the face
    const double tiny_h = DBL_MIN;
    if ((ur == 0.0) || ((ur < 0.0) && (hp <= tiny_h)) || ((ur > 0.0) && (hm <= tiny_h))) {
      uh = 0.0; cf = 0.0;
    } else if (ur < 0.0) {
      double hup = hp - areaT[q2 + st] * min_h;
      double hlos = gmax2(0.0, res[q + st]);
      if ((((hup - hlos) + ur) < 0.0) && ((0.5 * hup + ur) < 0.0)) {
        uh = gmin3(-0.5 * hup, -hup + hlos, 0.0);
        domore[fl] = 1;
      } else {
        uh = ur;
      }
      cf = -uh / hp;
    } else {
      double hup = hm - areaT[q2] * min_h;
      double hlos = gmax2(0.0, -res[q - st]);
      if ((((hup - hlos) - ur) < 0.0) && ((0.5 * hup - ur) < 0.0)) {
        uh = gmax3(0.5 * hup, hup - hlos, 0.0);
        domore[fl] = 1;
      } else {
        uh = ur;
      }
      cf = uh / hm;
    }
  }
  hh[q] = uh;
  if (cfl) cfl[q] = cf;      // (the streaming sweeps read 0.5*CFL and 1-2/3*CFL instead: the dense first trip skips this array)
  if (hcv) {      // per-face factors of the flux formulas, shared by every tracer (streaming kernels)
    hcv[q] = 0.5 * cf;                                   // 0.5*CFL          (:551,554 / :937,941)
    gv[q] = plm ? (1. - cf) : (1. - (2. / 3.) * cf);     // 1-CFL (:565,569) or 1-2/3*CFL
  }
}
template <int DIR>
__device__ __forceinline__ bool mass_face_ok(const Range &rg, int i, int j) {
  return (DIR == 0) ? (i >= rg.is - 1 && i <= rg.ie && j >= rg.js && j <= rg.je)
                    : (i >= rg.is && i <= rg.ie && j >= rg.js - 1 && j <= rg.je);
}

// dense form: one thread per face of the padded plane
template <int DIR>
__global__ void k_mass_flux(Geom g, Range rg, const double *__restrict__ res, const double *__restrict__ hprev,
                            const double *__restrict__ areaT, double min_h, const int *__restrict__ entry,
                            int *__restrict__ domore, double *__restrict__ hh, double *__restrict__ cfl,
                            double *__restrict__ hcv, double *__restrict__ gv, int plm) {
  const int ci = blockIdx.x * blockDim.x + threadIdx.x, rj = blockIdx.y, k0 = blockIdx.z;
  if (ci >= g.P) return;
  const int i = ci - g.coff, j = rj - g.roff;
  const long long q = gidx(g, ci, rj, k0), q2 = gidx(g, ci, rj, 0);
  const bool swept = mass_face_ok<DIR>(rg, i, j) && entry[k0 * g.R + rj];
  mass_flux_face<DIR>(g, q, q2, k0 * g.R + rj, swept, swept ? res[q] : 0.0, res, hprev, areaT, min_h, domore, hh, cfl, hcv, gv, plm);
}

// Residual trips: uhr / vhr are zero almost everywhere, so a 32-face tile (one warp) whose faces -- and the one
// face on either side in the sweep direction -- carry no residual transport has only neutral outputs
// (uhh = CFL = 0) that nothing reads: the list-driven tracer update (sparse.cu) touches only cells next to a
// non-zero face.  A persistent grid scans uhr / vhr, writes the outputs of the active tiles only, records per tile
// bit 0 "outputs written" and bit 1 "a cell of this tile may change" (a residual on the tile's own faces or on the
// face before them) and appends the tiles with bit 1 to tlist, which is all k_mass_update_list visits.
#define SCAN_UNROLL 4
template <int DIR>
__global__ void __launch_bounds__(256)
k_mass_flux_scan(Geom g, Range rg, const double *__restrict__ res, const double *__restrict__ hprev,
                 const double *__restrict__ areaT, double min_h, const int *__restrict__ entry,
                 int *__restrict__ domore, double *__restrict__ hh, double *__restrict__ cfl,
                 double *__restrict__ hcv, double *__restrict__ gv, int plm, int *__restrict__ tflags,
                 int *__restrict__ tlist, int *__restrict__ tcount) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int W = g.P >> 5;                                // warp tiles per row (P is a multiple of 64)
  const int nrow = g.R * g.nk;
  const int nwarp = gridDim.x * (blockDim.x >> 5);
  const int st = (DIR == 0) ? 1 : g.P;
  // one warp per row of a level at a time: the row's flags and ranges are decoded once, its tiles are streamed
  for (int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); row < nrow; row += nwarp) {
    const int k0 = row / g.R, rj = row - k0 * g.R, j = rj - g.roff;
    const bool row_ok = ((DIR == 0) ? (j >= rg.js && j <= rg.je) : (j >= rg.js - 1 && j <= rg.je)) && entry[row];
    bool lo_ok = false, hi_ok = false;                   // y: the face rows south / north of this one
    if (DIR == 1) {
      lo_ok = rj > 0 && (j - 1 >= rg.js - 1 && j - 1 <= rg.je) && entry[row - 1];
      hi_ok = rj + 1 < g.R && (j + 1 >= rg.js - 1 && j + 1 <= rg.je) && entry[row + 1];
    }
    const int t0 = row * W;
    if (!row_ok && !lo_ok && !hi_ok) {                   // nothing swept near this row: every tile is clean
      for (int wx = lane; wx < W; wx += 32) tflags[t0 + wx] = 0;
      continue;
    }
    const long long qrow = (long long)g.P * rj + g.plane * k0;
    const int i_lo = (DIR == 0) ? rg.is - 1 : rg.is;     // face columns of the sweep
    for (int wb = 0; wb < W; wb += SCAN_UNROLL) {
      double ur[SCAN_UNROLL], lo[SCAN_UNROLL], hi[SCAN_UNROLL];
#pragma unroll
      for (int u = 0; u < SCAN_UNROLL; u++) {            // the loads of the batch first
        ur[u] = 0.0; lo[u] = 0.0; hi[u] = 0.0;
        const int wx = wb + u;
        if (wx >= W) continue;
        const int ci = wx * 32 + lane, i = ci - g.coff;
        const long long q = qrow + ci;
        const bool col_ok = (i >= i_lo && i <= rg.ie);
        if (row_ok && col_ok) ur[u] = res[q];
        if (DIR == 0) {
          if (row_ok && lane == 0 && ci > 0 && i - 1 >= i_lo && i - 1 <= rg.ie) lo[u] = res[q - 1];
          if (row_ok && lane == 31 && ci + 1 < g.P && i + 1 >= i_lo && i + 1 <= rg.ie) hi[u] = res[q + 1];
        } else {
          if (lo_ok && col_ok) lo[u] = res[q - st];
          if (hi_ok && col_ok) hi[u] = res[q + st];
        }
      }
#pragma unroll
      for (int u = 0; u < SCAN_UNROLL; u++) {
        const int wx = wb + u;
        if (wx >= W) continue;                           // warp-uniform
        const bool nz = (ur[u] != 0.0), nz_lo = (lo[u] != 0.0), nz_hi = (hi[u] != 0.0);
        const int any_flux = __any_sync(FULL, nz || nz_lo || nz_hi);
        const int any_cell = __any_sync(FULL, nz || nz_lo);
        if (lane == 0) {
          tflags[t0 + wx] = (any_flux ? 1 : 0) | (any_cell ? 2 : 0);
          if (any_cell) tlist[atomicAdd(tcount, 1)] = t0 + wx;
        }
        if (!any_flux) continue;
        const int ci = wx * 32 + lane, i = ci - g.coff;
        const long long q = qrow + ci, q2 = (long long)g.P * rj + ci;
        if (DIR == 0 && lane == 0 && ci > 0 && !nz_lo) {
          // the face before the tile may belong to a skipped tile: its neutral values are read by this tile's first cell
          hh[q - 1] = 0.0; cfl[q - 1] = 0.0;
          if (hcv) { hcv[q - 1] = 0.0; gv[q - 1] = 1.0; }
        }
        const bool swept = row_ok && (i >= i_lo && i <= rg.ie);
        mass_flux_face<DIR>(g, q, q2, row, swept, ur[u], res, hprev, areaT, min_h, domore, hh, cfl, hcv, gv, plm);
      }
    }
  }
}
void launch_mass_flux(cudaStream_t s, int dir, const Geom &g, Range rg, const double *res, const double *hprev,
                      const double *areaT, double min_h, const int *entry, int *domore, double *hh, double *cfl,
                      double *hc, double *gq, int plm) {
  dim3 grid((g.P + TPB - 1) / TPB, g.R, g.nk);
  if (dir == 0) k_mass_flux<0><<<grid, TPB, 0, s>>>(g, rg, res, hprev, areaT, min_h, entry, domore, hh, cfl, hc, gq, plm);
  else k_mass_flux<1><<<grid, TPB, 0, s>>>(g, rg, res, hprev, areaT, min_h, entry, domore, hh, cfl, hc, gq, plm);
}
long long mass_tile_count(const Geom &g) { return (long long)(g.P / 32) * g.R * g.nk; }
// residual-trip form; tcount must be zero on entry
void launch_mass_flux_scan(cudaStream_t s, int dir, const Geom &g, Range rg, const double *res, const double *hprev,
                           const double *areaT, double min_h, const int *entry, int *domore, double *hh, double *cfl,
                           double *hc, double *gq, int plm, int *tflags, int *tlist, int *tcount, int nsm) {
  const int grid = nsm * 8;
  if (dir == 0) k_mass_flux_scan<0><<<grid, 256, 0, s>>>(g, rg, res, hprev, areaT, min_h, entry, domore, hh, cfl, hc, gq, plm, tflags, tlist, tcount);
  else k_mass_flux_scan<1><<<grid, 256, 0, s>>>(g, rg, res, hprev, areaT, min_h, entry, domore, hh, cfl, hc, gq, plm, tflags, tlist, tcount);
}

// ---------------------------------------------------------------------------
// Residual transport and cell volume update (:632-649 for x, :1021-1040 for y).
// ihnew == 0 encodes do_i == .false. (a true Ihnew is always > 0).
// ---------------------------------------------------------------------------
template <int DIR>
__device__ __forceinline__ void mass_update_cell(const Geom &g, const Range &rg, int ci, int rj, int k0,
                                                 double *__restrict__ res, double *__restrict__ hprev,
                                                 const double *__restrict__ hh, const double *__restrict__ areaT,
                                                 const double *__restrict__ neglect, double h_neglect,
                                                 const int *__restrict__ entry, const int *__restrict__ domore_k,
                                                 double *__restrict__ hlst, double *__restrict__ ihnew,
                                                 int *__restrict__ list, int *__restrict__ count, int cap,
                                                 int *__restrict__ nz = nullptr) {
  const int i = ci - g.coff, j = rj - g.roff;
  const long long q = gidx(g, ci, rj, k0), q2 = gidx(g, ci, rj, 0);
  const int st = (DIR == 0) ? 1 : g.P;
  const bool lev = domore_k[k0] > 0;
  const bool in_i = (i >= rg.is && i <= rg.ie), in_j = (j >= rg.js && j <= rg.je);
  bool face_ok, cell_ok;
  if (DIR == 0) {
    const bool row = lev && in_j && entry[k0 * g.R + rj];
    face_ok = row && (i >= rg.is - 1 && i <= rg.ie);
    cell_ok = row && in_i;
  } else {
    face_ok = lev && in_i && (j >= rg.js - 1 && j <= rg.je);
    cell_ok = lev && in_i && in_j;
  }
  const double uh = hh[q];
  if (face_ok) {
    double r = res[q] - uh;
    if (fabs(r) < neglect[q2]) r = 0.0;
    res[q] = r;
    // dense trip: count the faces that keep a residual -- twice that number estimates the cells the next trip lists
    // (the host only uses it to choose between the list-driven and the dense kernels: a hint, never a bound)
    if (nz && r != 0.0) atomicAdd(nz, 1);
  }
  double hl = 0.0, ih = 0.0;
  if (cell_ok) {
    const double uhm = hh[q - st];
    if ((uh != 0.0) || (uhm != 0.0)) {
      hl = hprev[q];
      double hn = hl - (uh - uhm);
      if (DIR == 1) hn = gmax2(hn, 0.0);
      hprev[q] = hn;
      const double hnegA = h_neglect * areaT[q2];
      if (hn <= 0.0) { ih = 0.0; }
      else if (hn < hnegA) { hl = hl + (hnegA - hn); ih = 1.0 / hnegA; }
      else { ih = 1.0 / hn; }
    }
  }
  hlst[q] = hl; ihnew[q] = ih;
  if (list && ih > 0.0) {      // residual iterations: remember the cells that change (sparse.cu)
    const int idx = atomicAdd(count, 1);
    if (idx < cap) list[idx] = (int)q;
  }
}
template <int DIR>
__global__ void k_mass_update(Geom g, Range rg, double *__restrict__ res, double *__restrict__ hprev,
                              const double *__restrict__ hh, const double *__restrict__ areaT,
                              const double *__restrict__ neglect, double h_neglect,
                              const int *__restrict__ entry, const int *__restrict__ domore_k,
                              double *__restrict__ hlst, double *__restrict__ ihnew,
                              int *__restrict__ list, int *__restrict__ count, int cap, int *__restrict__ nz) {
  const int ci = blockIdx.x * blockDim.x + threadIdx.x, rj = blockIdx.y, k0 = blockIdx.z;
  if (ci >= g.P) return;
  mass_update_cell<DIR>(g, rg, ci, rj, k0, res, hprev, hh, areaT, neglect, h_neglect, entry, domore_k, hlst, ihnew, list, count, cap, nz);
}
// residual trips: only the tiles k_mass_flux_scan listed (nothing changes in the others)
template <int DIR>
__global__ void __launch_bounds__(256)
k_mass_update_list(Geom g, Range rg, double *__restrict__ res, double *__restrict__ hprev,
                   const double *__restrict__ hh, const double *__restrict__ areaT,
                   const double *__restrict__ neglect, double h_neglect,
                   const int *__restrict__ entry, const int *__restrict__ domore_k,
                   double *__restrict__ hlst, double *__restrict__ ihnew,
                   int *__restrict__ list, int *__restrict__ count, int cap,
                   const int *__restrict__ tlist, const int *__restrict__ tcount) {
  const int lane = threadIdx.x & 31;
  const int W = g.P >> 5;
  const int n = *tcount;
  const int nwarp = gridDim.x * (blockDim.x >> 5);
  for (int e = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); e < n; e += nwarp) {
    const int t = tlist[e];
    const int wx = t % W, rk = t / W, rj = rk % g.R, k0 = rk / g.R;
    mass_update_cell<DIR>(g, rg, wx * 32 + lane, rj, k0, res, hprev, hh, areaT, neglect, h_neglect, entry, domore_k, hlst, ihnew, list, count, cap);
  }
}
void launch_mass_update(cudaStream_t s, int dir, const Geom &g, Range rg, double *res, double *hprev,
                        const double *hh, const double *areaT, const double *neglect, double h_neglect,
                        const int *entry, const int *domore_k, double *hlst, double *ihnew,
                        int *list, int *count, int cap, int *nz) {
  dim3 grid((g.P + TPB - 1) / TPB, g.R, g.nk);
  if (dir == 0) k_mass_update<0><<<grid, TPB, 0, s>>>(g, rg, res, hprev, hh, areaT, neglect, h_neglect, entry, domore_k, hlst, ihnew, list, count, cap, nz);
  else k_mass_update<1><<<grid, TPB, 0, s>>>(g, rg, res, hprev, hh, areaT, neglect, h_neglect, entry, domore_k, hlst, ihnew, list, count, cap, nz);
}
void launch_mass_update_list(cudaStream_t s, int dir, const Geom &g, Range rg, double *res, double *hprev,
                             const double *hh, const double *areaT, const double *neglect, double h_neglect,
                             const int *entry, const int *domore_k, double *hlst, double *ihnew,
                             int *list, int *count, int cap, const int *tlist, const int *tcount, int nsm) {
  const int grid = nsm * 4;
  if (dir == 0) k_mass_update_list<0><<<grid, 256, 0, s>>>(g, rg, res, hprev, hh, areaT, neglect, h_neglect, entry, domore_k, hlst, ihnew, list, count, cap, tlist, tcount);
  else k_mass_update_list<1><<<grid, 256, 0, s>>>(g, rg, res, hprev, hh, areaT, neglect, h_neglect, entry, domore_k, hlst, ihnew, list, count, cap, tlist, tcount);
}

// ---------------------------------------------------------------------------
// Tracer sweep (unfused path): every tracer of the registry in one launch,
// out of place (buf[src] -> buf[1-src]); a thread owns one cell and keeps the
// tracer-independent mass quantities in registers across the tracer loop.
//   fluxes  :517-577 / :902-963     update :652-662 / :1043-1047
//   ad_x/ad_y :665-667 / :1070-1076  advection_xy :671-676 / :1051-1056
//   underflow :683-687 / :1062-1066
// W = tracer stencil half width actually read (2 for PLM and PPM:H3, 3 for PPM-CW).
// ---------------------------------------------------------------------------
template <int SCHEME, int DIR>
__global__ void __launch_bounds__(TPB)
k_tracer_sweep(Geom g, Range rg, const TracerDev *__restrict__ tr, int ntr, int src,
               const double *__restrict__ hh, const double *__restrict__ cflv,
               const double *__restrict__ hlstv, const double *__restrict__ ihnewv,
               const double *__restrict__ mmv, const double *__restrict__ IareaT,
               const int *__restrict__ entry, const int *__restrict__ domore_k, double Idt) {
  int ci = blockIdx.x * blockDim.x + threadIdx.x, rj = blockIdx.y, k0 = blockIdx.z;
  if (ci >= g.P) return;
  const int i = ci - g.coff, j = rj - g.roff;
  if (i < g.isd || i > g.ied || j < g.jsd || j > g.jed) return;
  const long long q = gidx(g, ci, rj, k0), q2 = gidx(g, ci, rj, 0);
  const int st = (DIR == 0) ? 1 : g.P;
  constexpr int W = (SCHEME == SCHEME_PPM_CW) ? 3 : 2;
  const bool lev = domore_k[k0] > 0;
  const bool in_rng = (i >= rg.is && i <= rg.ie && j >= rg.js && j <= rg.je);
  const bool hi_ok = (DIR == 0) ? (i >= rg.is - 1 && i <= rg.ie && j >= rg.js && j <= rg.je)
                                : (i >= rg.is && i <= rg.ie && j >= rg.js - 1 && j <= rg.je);
  const bool act_hi = lev && hi_ok && entry[k0 * g.R + rj];
  const bool act_lo = (DIR == 0) ? act_hi : (lev && in_rng && entry[k0 * g.R + rj - 1]);
  const double ih = (lev && in_rng) ? ihnewv[q] : 0.0;
  const bool do_i = ih > 0.0;
  // do_i of the next cell (east / north); outside the sweep range it reads as .false.
  const bool nxt_in = (DIR == 0) ? (i + 1 >= rg.is && i + 1 <= rg.ie && j >= rg.js && j <= rg.je)
                                 : (i >= rg.is && i <= rg.ie && j + 1 >= rg.js && j + 1 <= rg.je);
  const bool do_nxt = lev && nxt_in && (ihnewv[q + st] > 0.0);
  const bool diag_hi = act_hi && (do_i || do_nxt);
  double uh_hi = 0.0, uh_lo = 0.0, cf_hi = 0.0, cf_lo = 0.0, hl = 0.0, mm[5] = {0, 0, 0, 0, 0}, iarea = 0.0;
  if (do_i || diag_hi) {
    uh_hi = hh[q]; cf_hi = cflv[q];
    uh_lo = hh[q - st]; cf_lo = cflv[q - st];
    hl = hlstv[q];
#pragma unroll
    for (int s = -2; s <= 2; s++) mm[s + 2] = mmv[q2 + s * st];
    iarea = IareaT[q2];
  }
  for (int m = 0; m < ntr; m++) {
    const double *__restrict__ tin = tr[m].buf[src];
    double *__restrict__ tout = tr[m].buf[1 - src];
    const double Tc = tin[q];
    double out = Tc;
    if (do_i || diag_hi) {
      double t[7];
#pragma unroll
      for (int s = -3; s <= 3; s++) t[s + 3] = (s >= -W && s <= W) ? tin[q + s * st] : 0.0;
      double f_hi = 0.0, f_lo = 0.0;
      if (act_hi) f_hi = face_flux<SCHEME>(&t[1], &mm[1], uh_hi, cf_hi);
      if (do_i) {
        if (act_lo) f_lo = face_flux<SCHEME>(&t[0], &mm[0], uh_lo, cf_lo);
        out = (Tc * hl - (f_hi - f_lo)) * ih;
        double *axy = tr[m].adv_xy;
        if (axy) axy[q] = axy[q] - (f_hi - f_lo) * Idt * iarea;
      }
      double *ad = (DIR == 0) ? tr[m].ad_x : tr[m].ad_y;
      if (ad && diag_hi) ad[q] = ad[q] + f_hi * Idt;
    }
    const double cu = tr[m].conc_underflow;
    if (lev && in_rng && cu > 0.0 && fabs(out) < cu) out = 0.0;
    tout[q] = out;
  }
}

template <int SCHEME>
static void launch_tracer_sweep_s(cudaStream_t s, int dir, const Geom &g, Range rg, const TracerDev *tr, int ntr,
                                  int src, const double *hh, const double *cfl, const double *hlst,
                                  const double *ihnew, const double *mm, const double *IareaT,
                                  const int *entry, const int *domore_k, double Idt) {
  dim3 grid((g.P + TPB - 1) / TPB, g.R, g.nk);
  if (dir == 0) k_tracer_sweep<SCHEME, 0><<<grid, TPB, 0, s>>>(g, rg, tr, ntr, src, hh, cfl, hlst, ihnew, mm, IareaT, entry, domore_k, Idt);
  else k_tracer_sweep<SCHEME, 1><<<grid, TPB, 0, s>>>(g, rg, tr, ntr, src, hh, cfl, hlst, ihnew, mm, IareaT, entry, domore_k, Idt);
}
void launch_tracer_sweep(cudaStream_t s, int scheme, int dir, const Geom &g, Range rg, const TracerDev *tr, int ntr,
                         int src, const double *hh, const double *cfl, const double *hlst, const double *ihnew,
                         const double *mm, const double *IareaT, const int *entry, const int *domore_k, double Idt) {
  if (scheme == SCHEME_PLM) launch_tracer_sweep_s<SCHEME_PLM>(s, dir, g, rg, tr, ntr, src, hh, cfl, hlst, ihnew, mm, IareaT, entry, domore_k, Idt);
  else if (scheme == SCHEME_PPM_CW) launch_tracer_sweep_s<SCHEME_PPM_CW>(s, dir, g, rg, tr, ntr, src, hh, cfl, hlst, ihnew, mm, IareaT, entry, domore_k, Idt);
  else launch_tracer_sweep_s<SCHEME_PPM_H3>(s, dir, g, rg, tr, ntr, src, hh, cfl, hlst, ihnew, mm, IareaT, entry, domore_k, Idt);
}

// ---------------------------------------------------------------------------
// Flux diagnostics of a sweep without the sweep (:665-667,671-676 / :1051-1056,1070-1076): the same face fluxes as
// k_tracer_sweep, from the tracers' values BEFORE the sweep, accumulated into ad_x | ad_y and advection_xy of the tracers
// that have them.  The streaming y march updates in place and carries no diagnostics; this runs in front of it for the
// handful of tracers a diag_table asks for, so that one T_ady no longer takes the whole registry off the fast path.
// ---------------------------------------------------------------------------
// One thread = one column of a FD_ROWS-row segment, marching north with the 7-point tracer window in registers, so
// that every tracer value is read once (plus the window's lead-in per segment) instead of once per stencil point.
#define FD_ROWS 64
template <int SCHEME>
__global__ void __launch_bounds__(TPB)
k_flux_diag_y(Geom g, Range rg, const TracerDev *__restrict__ tr, int ntr, int src,
              const double *__restrict__ hh, const double *__restrict__ cflv, const double *__restrict__ ihnewv,
              const double *__restrict__ mmv, const double *__restrict__ IareaT,
              const int *__restrict__ entry, const int *__restrict__ domore_k, double Idt) {
  const int ci = blockIdx.x * blockDim.x + threadIdx.x, k0 = blockIdx.z;
  if (ci >= g.P || domore_k[k0] <= 0) return;
  const int i = ci - g.coff;
  if (i < rg.is || i > rg.ie) return;                 // faces and cells of the sweep lie in columns is..ie
  constexpr int W = (SCHEME == SCHEME_PPM_CW) ? 3 : 2;
  const int P = g.P;
  // cell rows this thread visits: faces J in [js-1, je] are the high faces of cells j in [js-1, je]
  const int rlo = max(rg.js - 1 + g.roff, blockIdx.y * FD_ROWS), rhi = min(rg.je + g.roff, blockIdx.y * FD_ROWS + FD_ROWS - 1);
  if (rlo > rhi) return;
  const int *__restrict__ ent = entry + k0 * g.R;
  const long long col = (long long)ci + g.plane * k0;
  for (int m = 0; m < ntr; m++) {
    double *ad = tr[m].ad_y;
    double *axy = tr[m].adv_xy;
    if (!ad && !axy) continue;
    const double *__restrict__ tin = tr[m].buf[src] + col;
    double t[7];                                      // T(r-3 .. r+3) of the current row r
#pragma unroll
    for (int s = -3; s < 3; s++) { const int rr = rlo + s; t[s + 3] = (s >= -W && rr >= 0 && rr < g.R) ? tin[(long long)P * rr] : 0.0; }
    for (int r = rlo; r <= rhi; r++) {
      t[6] = (W == 3 && r + 3 < g.R) ? tin[(long long)P * (r + 3)] : 0.0;
      if (W == 2) t[5] = (r + 2 < g.R) ? tin[(long long)P * (r + 2)] : 0.0;
      const int j = r - g.roff;
      const long long q = col + (long long)P * r, q2 = (long long)ci + (long long)P * r;
      const bool in_rng = (j >= rg.js && j <= rg.je);
      const bool act_hi = (j >= rg.js - 1 && j <= rg.je) && ent[r];
      const bool act_lo = in_rng && ent[r - 1];
      const bool do_i = in_rng && ihnewv[q] > 0.0;
      const bool do_nxt = (j + 1 >= rg.js && j + 1 <= rg.je) && (ihnewv[q + P] > 0.0);
      const bool diag_hi = act_hi && (do_i || do_nxt);
      if (do_i || diag_hi) {
        double mm[5];
#pragma unroll
        for (int s = -2; s <= 2; s++) mm[s + 2] = mmv[q2 + (long long)s * P];
        double f_hi = 0.0, f_lo = 0.0;
        if (act_hi) f_hi = face_flux<SCHEME>(&t[1], &mm[1], hh[q], cflv[q]);
        if (do_i && axy) {
          if (act_lo) f_lo = face_flux<SCHEME>(&t[0], &mm[0], hh[q - P], cflv[q - P]);
          axy[q] = axy[q] - (f_hi - f_lo) * Idt * IareaT[q2];
        }
        if (ad && diag_hi) ad[q] = ad[q] + f_hi * Idt;
      }
#pragma unroll
      for (int s = 0; s < 6; s++) t[s] = t[s + 1];
    }
  }
}
void launch_flux_diag(cudaStream_t s, int scheme, int dir, const Geom &g, Range rg, const TracerDev *tr, int ntr, int src,
                      const double *hh, const double *cfl, const double *ihnew, const double *mm, const double *IareaT,
                      const int *entry, const int *domore_k, double Idt) {
  (void)dir;      // y only: the x sweep carries its diagnostics itself (k_x_fast<.., true>)
  dim3 grid((g.P + TPB - 1) / TPB, (g.R + FD_ROWS - 1) / FD_ROWS, g.nk);
  if (scheme == SCHEME_PLM) k_flux_diag_y<SCHEME_PLM><<<grid, TPB, 0, s>>>(g, rg, tr, ntr, src, hh, cfl, ihnew, mm, IareaT, entry, domore_k, Idt);
  else if (scheme == SCHEME_PPM_CW) k_flux_diag_y<SCHEME_PPM_CW><<<grid, TPB, 0, s>>>(g, rg, tr, ntr, src, hh, cfl, ihnew, mm, IareaT, entry, domore_k, Idt);
  else k_flux_diag_y<SCHEME_PPM_H3><<<grid, TPB, 0, s>>>(g, rg, tr, ntr, src, hh, cfl, ihnew, mm, IareaT, entry, domore_k, Idt);
}

// ---------------------------------------------------------------------------
// ad2d_x / ad2d_y (:691-699 / :1078-1084): vertical sums of the face fluxes in
// the reference's k order.  One thread per face column walks k = 1..nk and
// recomputes the flux from the sweep's (still intact) input buffer.
// ---------------------------------------------------------------------------
template <int SCHEME, int DIR>
__global__ void k_ad2d(Geom g, Range rg, const TracerDev *__restrict__ tr, int ntr, int src,
                       const double *__restrict__ hh, const double *__restrict__ cflv,
                       const double *__restrict__ ihnewv, const double *__restrict__ mmv,
                       const int *__restrict__ entry, const int *__restrict__ domore_k, double Idt,
                       const int *__restrict__ tflags) {
  // tflags (list-driven residual trips, nullable): Ihnew of a cell is current only where the mass update visited its
  // 32-column tile (bit 1); elsewhere no cell changes in this sweep
  int ci = blockIdx.x * blockDim.x + threadIdx.x, rj = blockIdx.y;
  if (ci >= g.P) return;
  const int i = ci - g.coff, j = rj - g.roff;
  const bool hi_ok = (DIR == 0) ? (i >= rg.is - 1 && i <= rg.ie && j >= rg.js && j <= rg.je)
                                : (i >= rg.is && i <= rg.ie && j >= rg.js - 1 && j <= rg.je);
  if (!hi_ok) return;
  const int st = (DIR == 0) ? 1 : g.P;
  constexpr int W = (SCHEME == SCHEME_PPM_CW) ? 3 : 2;
  const bool cur_in = (i >= rg.is && i <= rg.ie && j >= rg.js && j <= rg.je);
  const bool nxt_in = (DIR == 0) ? (i + 1 <= rg.ie) : (j + 1 <= rg.je);
  const long long q2 = gidx(g, ci, rj, 0);
  double mm[4];
#pragma unroll
  for (int s = -1; s <= 2; s++) mm[s + 1] = mmv[q2 + s * st];
  for (int m = 0; m < ntr; m++) {
    double *a2 = (DIR == 0) ? tr[m].ad2d_x : tr[m].ad2d_y;
    if (!a2) continue;
    const double *__restrict__ tin = tr[m].buf[src];
    double acc = a2[q2];
    for (int k0 = 0; k0 < g.nk; k0++) {
      if (domore_k[k0] <= 0 || !entry[k0 * g.R + rj]) continue;
      const long long q = q2 + g.plane * k0;
      bool vc = true, vn = true;
      if (tflags) {
        const int W5 = g.P >> 5;
        vc = (tflags[(k0 * g.R + rj) * W5 + (ci >> 5)] & 2) != 0;
        vn = (DIR == 0) ? ((ci + 1 < g.P) && (tflags[(k0 * g.R + rj) * W5 + ((ci + 1) >> 5)] & 2) != 0)
                        : ((rj + 1 < g.R) && (tflags[(k0 * g.R + rj + 1) * W5 + (ci >> 5)] & 2) != 0);
      }
      const bool d = (cur_in && vc && ihnewv[q] > 0.0) || (nxt_in && vn && ihnewv[q + st] > 0.0);
      if (!d) continue;
      double t[6];
#pragma unroll
      for (int s = -2; s <= 3; s++) t[s + 2] = (s >= 1 - W && s <= W) ? tin[q + s * st] : 0.0;
      acc = acc + face_flux<SCHEME>(t, mm, hh[q], cflv[q]) * Idt;
    }
    a2[q2] = acc;
  }
}
void launch_ad2d(cudaStream_t s, int scheme, int dir, const Geom &g, Range rg, const TracerDev *tr, int ntr, int src,
                 const double *hh, const double *cfl, const double *ihnew, const double *mm, const int *entry,
                 const int *domore_k, double Idt, const int *tflags) {
  dim3 grid((g.P + TPB - 1) / TPB, g.R, 1);
#define AD2D(S, D) k_ad2d<S, D><<<grid, TPB, 0, s>>>(g, rg, tr, ntr, src, hh, cfl, ihnew, mm, entry, domore_k, Idt, tflags)
  if (scheme == SCHEME_PLM) { if (dir == 0) AD2D(SCHEME_PLM, 0); else AD2D(SCHEME_PLM, 1); }
  else if (scheme == SCHEME_PPM_CW) { if (dir == 0) AD2D(SCHEME_PPM_CW, 0); else AD2D(SCHEME_PPM_CW, 1); }
  else { if (dir == 0) AD2D(SCHEME_PPM_H3, 0); else AD2D(SCHEME_PPM_H3, 1); }
#undef AD2D
}

// ---------------------------------------------------------------------------
// C1: group pass on one device (:206).  The plan precomputes, per stagger, a
// list of (dst, src, sign) plane offsets that encodes closed / cyclic / north
// fold topology; one launch fills the halos of every listed field and level.
// Sources are computational-domain points, destinations halo points: disjoint.
// ---------------------------------------------------------------------------
__global__ void k_halo_apply(long long plane, int nk, const int *__restrict__ dst, const int *__restrict__ srcv,
                             const signed char *__restrict__ sgn, int n, double *const *__restrict__ fields,
                             int src_sel /* -1: fields are plain pointers */) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int k0 = blockIdx.y;
  double *f = fields[blockIdx.z];
  const long long o = plane * k0;
  double v = f[o + srcv[t]];
  f[o + dst[t]] = (sgn[t] < 0) ? -v : v;
  (void)nk; (void)src_sel;
}
void launch_halo_apply(cudaStream_t s, long long plane, int nk, const int *dst, const int *src,
                       const signed char *sgn, int n, double *const *fields, int nfields) {
  if (n <= 0 || nfields <= 0) return;
  dim3 grid((n + TPB - 1) / TPB, nk, nfields);
  k_halo_apply<<<grid, TPB, 0, s>>>(plane, nk, dst, src, sgn, n, fields, -1);
}

// Pack / unpack of halo strips for the multi-GPU exchange: buffer layout is
// [field][level][point] so that one message per neighbour carries all fields.
__global__ void k_halo_pack(long long plane, int nk, const int *__restrict__ srcv, const signed char *__restrict__ sgn,
                            int n, double *const *__restrict__ fields, double *__restrict__ buf) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int k0 = blockIdx.y, f = blockIdx.z;
  double v = fields[f][plane * k0 + srcv[t]];
  buf[((long long)f * nk + k0) * n + t] = (sgn[t] < 0) ? -v : v;
}
__global__ void k_halo_unpack(long long plane, int nk, const int *__restrict__ dstv, int n,
                              double *const *__restrict__ fields, const double *__restrict__ buf) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int k0 = blockIdx.y, f = blockIdx.z;
  fields[f][plane * k0 + dstv[t]] = buf[((long long)f * nk + k0) * n + t];
}
// Peer-pull form of pack + send + recv + unpack: every halo point is loaded straight from the owner's
// field in the neighbouring GPU's HBM (CUDA IPC mapping, NVLink load) and stored in place.
__global__ void k_halo_pull(long long plane, long long peer_plane, int nk, const int *__restrict__ dstv,
                            const int *__restrict__ srcv, const signed char *__restrict__ sgn, int n,
                            double *const *__restrict__ fields, double *const *__restrict__ peer_fields) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n) return;
  const int k0 = blockIdx.y, f = blockIdx.z;
  const double v = __ldcv(peer_fields[f] + peer_plane * k0 + srcv[t]);
  fields[f][plane * k0 + dstv[t]] = (sgn[t] < 0) ? -v : v;
}
void launch_halo_pull(cudaStream_t s, long long plane, long long peer_plane, int nk, const int *dst, const int *src,
                      const signed char *sgn, int n, double *const *fields, double *const *peer_fields, int nfields) {
  if (n <= 0 || nfields <= 0) return;
  dim3 grid((n + TPB - 1) / TPB, nk, nfields);
  k_halo_pull<<<grid, TPB, 0, s>>>(plane, peer_plane, nk, dst, src, sgn, n, fields, peer_fields);
}
// Pair-list form: entry t moves the two consecutive points dst2[t], dst2[t]+1 <- src2[t], src2[t]+1 (both even: 16-byte
// aligned) of four levels per thread, all loads issued before the first store: 64 bytes in flight per thread over NVLink.
// tile2[t] = the source points' 32-column tile in the OWNER's plane; the map (fetched into local memory first, so that
// the pull makes one NVLink round trip) says whether any cell of that tile changed since the last pass: after a
// list-driven residual trip nearly all of the strip is skipped.
#define PULL_LEV 4
__global__ void __launch_bounds__(128)
k_halo_pull2(long long plane, long long peer_plane, int nk, const int *__restrict__ dst2, const int *__restrict__ src2,
             const signed char *__restrict__ sgn2, const int *__restrict__ tile2, int n2,
             double *const *__restrict__ fields, double *const *__restrict__ peer_fields,
             const unsigned char *__restrict__ dirty, int tiles_plane) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n2) return;
  const int f = blockIdx.z, kb = blockIdx.y * PULL_LEV;
  const int d = dst2[t], sidx = src2[t];
  const bool neg = sgn2[t] < 0;
  bool need[PULL_LEV];
#pragma unroll
  for (int u = 0; u < PULL_LEV; u++) need[u] = (kb + u < nk);
  if (dirty && dirty[(long long)tiles_plane * nk] == 0) {
    const int tl = tile2[t];
#pragma unroll
    for (int u = 0; u < PULL_LEV; u++) if (need[u]) need[u] = dirty[(long long)tiles_plane * (kb + u) + tl] != 0;
  }
  const double *pf = peer_fields[f];
  double2 v[PULL_LEV];
#pragma unroll
  for (int u = 0; u < PULL_LEV; u++)
    if (need[u]) v[u] = __ldcv(reinterpret_cast<const double2 *>(pf + peer_plane * (kb + u) + sidx));
  double *mf = fields[f];
#pragma unroll
  for (int u = 0; u < PULL_LEV; u++)
    if (need[u]) {
      double2 w = v[u];
      if (neg) { w.x = -w.x; w.y = -w.y; }
      *reinterpret_cast<double2 *>(mf + plane * (kb + u) + d) = w;
    }
}
void launch_halo_pull2(cudaStream_t s, long long plane, long long peer_plane, int nk, const int *dst2, const int *src2,
                       const signed char *sgn2, const int *tile2, int n2, double *const *fields,
                       double *const *peer_fields, int nfields, const unsigned char *dirty, int peer_tiles_plane) {
  if (n2 <= 0 || nfields <= 0) return;
  dim3 grid((n2 + 127) / 128, (nk + PULL_LEV - 1) / PULL_LEV, nfields);
  k_halo_pull2<<<grid, 128, 0, s>>>(plane, peer_plane, nk, dst2, src2, sgn2, tile2, n2, fields, peer_fields, dirty, peer_tiles_plane);
}
__global__ void k_fetch_bytes(unsigned char *__restrict__ dst, const unsigned char *__restrict__ src, long long n16) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n16) reinterpret_cast<int4 *>(dst)[t] = __ldcv(reinterpret_cast<const int4 *>(src) + t);
}
void launch_fetch_bytes(cudaStream_t s, unsigned char *dst, const unsigned char *src, long long bytes) {
  const long long n16 = bytes / 16;      // (the maps are padded to a multiple of 16 bytes)
  if (n16 <= 0) return;
  k_fetch_bytes<<<(unsigned)((n16 + 255) / 256), 256, 0, s>>>(dst, src, n16);
}
void launch_halo_pack(cudaStream_t s, long long plane, int nk, const int *src, const signed char *sgn, int n,
                      double *const *fields, int nfields, double *buf) {
  if (n <= 0 || nfields <= 0) return;
  dim3 grid((n + TPB - 1) / TPB, nk, nfields);
  k_halo_pack<<<grid, TPB, 0, s>>>(plane, nk, src, sgn, n, fields, buf);
}
void launch_halo_unpack(cudaStream_t s, long long plane, int nk, const int *dst, int n,
                        double *const *fields, int nfields, const double *buf) {
  if (n <= 0 || nfields <= 0) return;
  dim3 grid((n + TPB - 1) / TPB, nk, nfields);
  k_halo_unpack<<<grid, TPB, 0, s>>>(plane, nk, dst, n, fields, buf);
}

__global__ void k_fill(double *p, long long n, double v) {
  long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) p[t] = v;
}
void launch_fill(cudaStream_t s, double *p, long long n, double v) {
  if (n <= 0) return;
  k_fill<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(p, n, v);
}

__global__ void k_fill_int(int *p, int n, int v) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) p[t] = v;
}
// Set on the device: a host-to-device copy here would queue behind the bulk tracer uploads of the MIRROR pipeline.
void launch_fill_int(cudaStream_t s, int *p, int n, int v) {
  if (n <= 0) return;
  k_fill_int<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(p, n, v);
}

// Publish n ints to host-mapped pinned memory: the convergence word and the active-cell count reach the
// host as posted PCIe writes, not through a copy engine busy with a chunk of tracers.
__global__ void k_publish(volatile int *host, const int *dev, int n) {
  if (threadIdx.x < n) host[threadIdx.x] = dev[threadIdx.x];
  __threadfence_system();
}
void launch_publish(cudaStream_t s, int *host_mapped, const int *dev, int n) {
  k_publish<<<1, 32, 0, s>>>(host_mapped, dev, n);
}
