// The iteration loop of offline_advection_ale (/root/reference/src/tracer/MOM_offline_main.F90:254-312, 320-411,
// 414-415) on the plan's resident fields: the element-wise statements around the advect_tracer call and
// remaining_transport_sum (:615-650) with its reproducing sum (src/framework/MOM_coms.F90:119-213).  The loop itself
// (plan.cu: mom6adv_offline_advect) keeps the reference's names:
//   h_pre / h_new -> the H_END field (the two are equal whenever either is read: :324 copies one into the other)
//   h_vol         -> the VOL_PREV field (advect_tracer's vol_prev argument, updated in place, :336)
//   uhtr_sub, vhtr_sub -> the UHTR / VHTR fields (advect_tracer's transports)
//   uhtr, vhtr    -> the UHR / VHR fields (advect_tracer's uhr_out / vhr_out, :336)
#include "internal.cuh"
#include "kernels.h"

#define OFF_TX 64
#define OFF_TY 4

// :270 + :304-306: h_new = 0 outside the computational domain (inside it is h_pre)
__global__ void __launch_bounds__(OFF_TX * OFF_TY)
k_off_init(Geom g, double *__restrict__ h) {
  const int col = blockIdx.x * OFF_TX + threadIdx.x, row = blockIdx.y * OFF_TY + threadIdx.y, k = blockIdx.z;
  if (col >= g.P || row >= g.R) return;
  const int i = col - g.coff, j = row - g.roff;
  if (i >= g.isc && i <= g.iec && j >= g.jsc && j <= g.jec) return;
  h[(long long)row * g.P + col + g.plane * k] = 0.0;
}

// :322-325  h_vol = h_new * areaT   (h_pre = h_new is the same field)
__global__ void __launch_bounds__(OFF_TX * OFF_TY)
k_off_vol(Geom g, const double *__restrict__ h, const double *__restrict__ areaT, double *__restrict__ vol) {
  const int i = g.isc + blockIdx.x * OFF_TX + threadIdx.x, j = g.jsc + blockIdx.y * OFF_TY + threadIdx.y, k = blockIdx.z;
  if (i > g.iec || j > g.jec) return;
  const long long q2 = (long long)(j + g.roff) * g.P + (i + g.coff), q = q2 + g.plane * k;
  vol[q] = h[q] * areaT[q2];
}

// :342-344  h_new = h_vol * IareaT ;  :382-385  uhtr_sub(I,j,k) = uhtr(I,j,k) ; vhtr_sub(i,J,k) = vhtr(i,J,k)  (i=is,ie ; j=js,je)
__global__ void __launch_bounds__(OFF_TX * OFF_TY)
k_off_post(Geom g, const double *__restrict__ vol, const double *__restrict__ IareaT, double *__restrict__ h,
           const double *__restrict__ uhr, const double *__restrict__ vhr, double *__restrict__ usub, double *__restrict__ vsub) {
  const int i = g.isc + blockIdx.x * OFF_TX + threadIdx.x, j = g.jsc + blockIdx.y * OFF_TY + threadIdx.y, k = blockIdx.z;
  if (i > g.iec || j > g.jec) return;
  const long long q2 = (long long)(j + g.roff) * g.P + (i + g.coff), q = q2 + g.plane * k;
  h[q] = vol[q] * IareaT[q2];
  usub[q] = uhr[q];
  vsub[q] = vhr[q];
}

// After the group pass of (hprev = h_new, uhr, vhr): everything outside the copies of :382-385 (i = is..ie, j = js..je)
// goes to h_new, uhtr_sub, vhtr_sub.  That includes the symmetric-memory boundary faces is-1 / js-1 where they have an
// owner: pass_vector delivers the neighbouring PE's (or the periodic image's) value of that same face, which is the
// value this PE computed for it itself -- uhr / vhr hold it.  On a closed edge of the global domain the face has no
// owner and keeps its value (keep_w / keep_s).
__global__ void __launch_bounds__(OFF_TX * OFF_TY)
k_off_halos(Geom g, int keep_w, int keep_s, const double *__restrict__ hprev, const double *__restrict__ uhr,
            const double *__restrict__ vhr, double *__restrict__ h, double *__restrict__ usub, double *__restrict__ vsub) {
  const int col = blockIdx.x * OFF_TX + threadIdx.x, row = blockIdx.y * OFF_TY + threadIdx.y, k = blockIdx.z;
  if (col >= g.P || row >= g.R) return;
  const int i = col - g.coff, j = row - g.roff;
  const long long q = (long long)row * g.P + col + g.plane * k;
  const bool iin = (i >= g.isc && i <= g.iec), jin = (j >= g.jsc && j <= g.jec);
  if (iin && jin) return;
  if (i >= g.isd && i <= g.ied && j >= g.jsd && j <= g.jed) h[q] = hprev[q];
  if (i >= g.IsdB && i <= g.ied && j >= g.jsd && j <= g.jed && !(keep_w && i == g.isc - 1 && jin)) usub[q] = uhr[q];
  if (i >= g.isd && i <= g.ied && j >= g.JsdB && j <= g.jed && !(keep_s && j == g.jsc - 1 && iin)) vsub[q] = vhr[q];
}

// remaining_transport_sum, :638-644, and the conversion of each column's value to extended-fixed-point integers
// (MOM_coms.F90:597-626 increment_ints_faster).  The integer sums are order-independent, so a block sums its columns'
// integers, carries exactly (every limb below 2**46 again) and adds them to six global accumulators; the host finishes
// with regularize_ints and ints_to_real (:653-691, :550-559).  acc[6] must be zero at launch.
__global__ void __launch_bounds__(256)
k_off_residual(Geom g, const double *__restrict__ u, const double *__restrict__ v, const double *__restrict__ h,
               const double *__restrict__ areaT, double Hsub, double HL2_to_kg_scale, unsigned long long *__restrict__ acc) {
  const int ni = g.iec - g.isc + 1, nj = g.jec - g.jsc + 1;
  const long long c = (long long)blockIdx.x * 256 + threadIdx.x;
  long long ints[6] = {0, 0, 0, 0, 0, 0};
  if (c < (long long)ni * nj) {
    const int i = g.isc + (int)(c % ni), j = g.jsc + (int)(c / ni);
    const long long q2 = (long long)(j + g.roff) * g.P + (i + g.coff);
    const double a2 = areaT[q2];
    double col = 0.0;                                                                  // :638
    for (int k = 0; k < g.nk; k++) {                                                   // :639
      const long long q = q2 + g.plane * k;
      const double trans_cell = (fabs(u[q - 1]) + fabs(u[q])) + (fabs(v[q - g.P]) + fabs(v[q]));     // :640-641
      if (trans_cell > gmax2(1.0e-16 * h[q], Hsub) * a2) col = col + HL2_to_kg_scale * trans_cell;   // :642-643
    }
    // col >= 0: sgn = 1
    const double r_prec = 70368744177664.0;      // 2**46
    const double pr[6] = {r_prec * r_prec, r_prec, 1.0, 1.0 / r_prec, 1.0 / (r_prec * r_prec), 1.0 / (r_prec * r_prec * r_prec)};
    const double I_pr[6] = {1.0 / (r_prec * r_prec), 1.0 / r_prec, 1.0, r_prec, r_prec * r_prec, r_prec * r_prec * r_prec};
    double rs = fabs(col);
#pragma unroll
    for (int n = 0; n < 6; n++) {
      const long long ival = (long long)(rs * I_pr[n]);
      rs = rs - (double)ival * pr[n];
      ints[n] = ival;
    }
  }
  __shared__ long long ws[8][6];
#pragma unroll
  for (int n = 0; n < 6; n++) {
    long long x = ints[n];
    for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5][n] = x;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    long long s[6];
    for (int n = 0; n < 6; n++) { s[n] = 0; for (int w = 0; w < 8; w++) s[n] += ws[w][n]; }
    const long long prec = 1LL << 46;
    for (int n = 5; n >= 1; n--) { const long long cy = s[n] >> 46; s[n] -= cy * prec; s[n - 1] += cy; }      // exact carry (s >= 0)
    for (int n = 0; n < 6; n++) if (s[n]) atomicAdd(acc + n, (unsigned long long)s[n]);
  }
}

void launch_off_init(cudaStream_t s, const Geom &g, double *h) {
  dim3 grid((g.P + OFF_TX - 1) / OFF_TX, (g.R + OFF_TY - 1) / OFF_TY, g.nk), block(OFF_TX, OFF_TY);
  k_off_init<<<grid, block, 0, s>>>(g, h);
}
void launch_off_halos(cudaStream_t s, const Geom &g, bool keep_w, bool keep_s, const double *hprev, const double *uhr,
                      const double *vhr, double *h, double *usub, double *vsub) {
  dim3 grid((g.P + OFF_TX - 1) / OFF_TX, (g.R + OFF_TY - 1) / OFF_TY, g.nk), block(OFF_TX, OFF_TY);
  k_off_halos<<<grid, block, 0, s>>>(g, keep_w ? 1 : 0, keep_s ? 1 : 0, hprev, uhr, vhr, h, usub, vsub);
}
void launch_off_vol(cudaStream_t s, const Geom &g, const double *h, const double *areaT, double *vol) {
  dim3 grid((g.iec - g.isc + OFF_TX) / OFF_TX, (g.jec - g.jsc + OFF_TY) / OFF_TY, g.nk), block(OFF_TX, OFF_TY);
  k_off_vol<<<grid, block, 0, s>>>(g, h, areaT, vol);
}
void launch_off_post(cudaStream_t s, const Geom &g, const double *vol, const double *IareaT, double *h, const double *uhr,
                     const double *vhr, double *usub, double *vsub) {
  dim3 grid((g.iec - g.isc + OFF_TX) / OFF_TX, (g.jec - g.jsc + OFF_TY) / OFF_TY, g.nk), block(OFF_TX, OFF_TY);
  k_off_post<<<grid, block, 0, s>>>(g, vol, IareaT, h, uhr, vhr, usub, vsub);
}
void launch_off_residual(cudaStream_t s, const Geom &g, const double *u, const double *v, const double *h,
                         const double *areaT, double Hsub, double HL2_to_kg_scale, unsigned long long *acc) {
  const long long ncol = (long long)(g.iec - g.isc + 1) * (g.jec - g.jsc + 1);
  k_off_residual<<<(unsigned)((ncol + 255) / 256), 256, 0, s>>>(g, u, v, h, areaT, Hsub, HL2_to_kg_scale, acc);
}
