// The along-layer diffusion iteration of tracer_hordiff
// (/root/reference/src/tracer/MOM_tracer_hor_diff.F90:539-607), on the plan's resident tracers.
//
//   do itt=1,num_itts
//     do_group_pass(CS%pass_t)                                  :540   (plan.cu: group pass, tracers only)
//     per level k: scale (:543-550), Coef_y (:552-555), Coef_x (:557-562), Ihdxdy (:563-565)
//     per tracer:  dTr (:569-574), t = t + dTr (:592-594)
//     conc_underflow (:600-605)
//
// h, khdt_x, khdt_y and scale do not change between iterations, so neither do Coef_x, Coef_y and Ihdxdy: the
// coefficient kernel runs once per call into three scratch fields and every iteration re-reads them (three loads per
// cell shared by all tracers, against five divisions per cell per iteration).  The reference updates a level in place
// after computing dTr for the whole level; here an iteration reads the live ping-pong buffer and writes the idle one,
// which is the same data flow.  Cells outside the computational domain and the levels :544-550 skip are copied through
// so that the new live buffer equals the reference's array everywhere (halos as the last pass left them).
//
// Arithmetic is the reference's expression order with FMA contraction off (build flags), IEEE division.
// HBM bytes per cell and iteration: 16 per tracer (read + write) + 24 (the three coefficient fields).
#include "internal.cuh"
#include "kernels.h"

__device__ __forceinline__ bool hd_level(const HdParams &p, int k1, double &scale) {   // k1: 1-based level, :543-550
  scale = p.I_numitts;
  if (p.diffuse_ml) {
    if (k1 <= p.nkml) {
      if (p.ml_scale <= 0.0) return false;
      scale = p.I_numitts * p.ml_scale;
    }
    if ((k1 > p.nkml) && (k1 <= p.nk_rho_varies)) return false;
  }
  return true;
}

#define HD_TX 64
#define HD_TY 4

// Coef_x (faces is-1..ie, rows js..je), Coef_y (columns is..ie, faces js-1..je), Ihdxdy (cells): :552-566
__global__ void __launch_bounds__(HD_TX * HD_TY)
k_hd_coef(Geom g, HdParams p, const double *__restrict__ h, const double *__restrict__ khdt_x,
          const double *__restrict__ khdt_y, const double *__restrict__ IareaT,
          double *__restrict__ cx, double *__restrict__ cy, double *__restrict__ ihd) {
  const int col = blockIdx.x * HD_TX + threadIdx.x, row = blockIdx.y * HD_TY + threadIdx.y, k = blockIdx.z;
  if (col >= g.P || row >= g.R) return;
  double scale;
  if (!hd_level(p, k + 1, scale)) return;
  const int i = col - g.coff, j = row - g.roff;
  const long long q2 = (long long)row * g.P + col, q = q2 + g.plane * k;
  const bool jin = (j >= g.jsc && j <= g.jec), iin = (i >= g.isc && i <= g.iec);
  if (iin && j >= g.jsc - 1 && j <= g.jec) {
    const double h0 = h[q], h1 = h[q + g.P];
    cy[q] = ((scale * khdt_y[q2]) * 2.0 * (h0 * h1)) / (h0 + h1 + p.h_neglect);
  }
  if (jin && i >= g.isc - 1 && i <= g.iec) {
    const double h0 = h[q], h1 = h[q + 1];
    cx[q] = ((scale * khdt_x[q2]) * 2.0 * (h0 * h1)) / (h0 + h1 + p.h_neglect);
  }
  if (iin && jin) ihd[q] = IareaT[q2] / (h[q] + p.h_neglect);
}

// One iteration for the tracers of tt: src (live) -> dst (idle).  A block owns a W x H tile of one level; a thread owns
// one cell and keeps its five coefficients in registers across the tracer loop.  The tracers' tiles (with their
// one-cell frame) are staged through shared memory by asynchronous 16-byte copies, NS tracers deep: every value
// crosses L2 -> SM once per tile instead of once per stencil arm, and the loads of the next NS-1 tracers are in flight
// while this one is computed, without holding registers.  (Measured against thread-per-cell kernels reading the five
// stencil points through L1, one or two cells per thread, U = 1..8 tracers unrolled: profiles/r02_hordiff.md.)
__device__ __forceinline__ void hd_cp_async16(double *smem_dst, const double *gsrc) {
  const unsigned sa = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void hd_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void hd_cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int W, int H, int NS>
__global__ void __launch_bounds__(W * H)
k_hd_tile(Geom g, HdParams p, const __grid_constant__ TrTable tt, const double *__restrict__ cx,
          const double *__restrict__ cy, const double *__restrict__ ihd) {
  constexpr int SP = W + 4, SROWS = H + 2, NCH = SROWS * (SP / 2), STAGE = SROWS * SP;
  static_assert(NCH <= W * H, "one staging chunk per thread");
  extern __shared__ __align__(16) double hd_sm[];
  const int tid = threadIdx.y * W + threadIdx.x;
  const int col0 = ((g.isd + g.coff) & ~1) + blockIdx.x * W, row0 = blockIdx.y * H, k = blockIdx.z;
  const int col = col0 + threadIdx.x, row = row0 + threadIdx.y;
  const int i = col - g.coff, j = row - g.roff;
  const bool live = (i >= g.isd && i <= g.ied && row < g.R);
  const long long q = (long long)row * g.P + col + g.plane * k;
  const bool inside = live && i >= g.isc && i <= g.iec && j >= g.jsc && j <= g.jec;
  double scale;
  if (!hd_level(p, k + 1, scale)) {      // uniform over the block (k = blockIdx.z)
    if (live)
      for (int m = 0; m < tt.n; m++) {
        double v = tt.src[m][q];
        if (inside && tt.cu[m] > 0.0 && fabs(v) < tt.cu[m]) v = 0.0;
        tt.dst[m][q] = v;
      }
    return;
  }
  // this thread's 16-byte piece of a stage: rows row0-1 .. row0+H, columns col0-2 .. col0+W+1
  const int crow = tid / (SP / 2), cc = tid % (SP / 2);
  const int grow = row0 - 1 + crow, gcol = col0 - 2 + 2 * cc;
  const bool chunk_ok = tid < NCH && grow >= 0 && grow < g.R && gcol >= 0 && gcol + 1 < g.P;
  const long long gq = (long long)grow * g.P + gcol + g.plane * k;
  const int so = crow * SP + 2 * cc;
  double cxm = 0.0, cxp = 0.0, cym = 0.0, cyp = 0.0, ih = 0.0;
  if (inside) { cxm = cx[q - 1]; cxp = cx[q]; cym = cy[q - g.P]; cyp = cy[q]; ih = ihd[q]; }
#pragma unroll
  for (int m = 0; m < NS - 1; m++) {
    if (m < tt.n && chunk_ok) hd_cp_async16(hd_sm + m * STAGE + so, tt.src[m] + gq);
    hd_cp_async_commit();
  }
  const int sc = (threadIdx.y + 1) * SP + threadIdx.x + 2;
  for (int m = 0; m < tt.n; m++) {
    hd_cp_async_wait<NS - 2>();          // this thread's piece of tracer m has landed ...
    __syncthreads();                     // ... and everybody's; everybody is also done with the stage refilled next
    const int mn = m + NS - 1;
    if (mn < tt.n && chunk_ok) hd_cp_async16(hd_sm + (mn % NS) * STAGE + so, tt.src[mn] + gq);
    hd_cp_async_commit();
    if (live) {
      const double *st = hd_sm + (m % NS) * STAGE + sc;
      const double c = st[0];
      double v = c;
      if (inside) {
        const double w = st[-1], e = st[1], sv = st[-SP], nv = st[SP];
        const double dTr = ih * ((cxm * (w - c) - cxp * (c - e)) + (cym * (sv - c) - cyp * (c - nv)));     // :569-574
        v = c + dTr;                                                                                       // :592-594
        const double cu = tt.cu[m];
        if (cu > 0.0 && fabs(v) < cu) v = 0.0;                                                             // :600-605
      }
      tt.dst[m][q] = v;
    }
  }
}

// ---- the diffusive-flux diagnostics, :576-591 (only launched when a tracer has one associated) ----------------------
// One iteration's contribution, from the tracer values BEFORE the update (the live buffers) and the coefficient fields:
//   df_x(I,j,k) += Coef_x(I,j) * (t(i,j,k) - t(i+1,j,k)) * Idt      j = js..je, I = G%IscB..G%IecB      (:576-579)
//   df_y(i,J,k) += Coef_y(i,J) * (t(i,j,k) - t(i,j+1,k)) * Idt      J = G%JscB..G%JecB, i = is..ie      (:580-583)
// G%IscB = isc-1 in symmetric memory and isc otherwise (MOM_hor_index.F90:87-96); levels the call skips (:544-550) get
// nothing.  Thread per point, tracer loop inside; every (point, tracer) is touched by one thread only.
__global__ void __launch_bounds__(HD_TX * HD_TY)
k_hd_df3(Geom g, HdParams p, const __grid_constant__ TrTable tt, const __grid_constant__ DfTable df,
         const double *__restrict__ cx, const double *__restrict__ cy, double Idt, int IscB, int JscB) {
  const int col = blockIdx.x * HD_TX + threadIdx.x, row = blockIdx.y * HD_TY + threadIdx.y, k = blockIdx.z;
  if (col >= g.P || row >= g.R) return;
  double scale;
  if (!hd_level(p, k + 1, scale)) return;
  const int i = col - g.coff, j = row - g.roff;
  const long long q = (long long)row * g.P + col + g.plane * k;
  const bool fx = (j >= g.jsc && j <= g.jec && i >= IscB && i <= g.iec);
  const bool fy = (i >= g.isc && i <= g.iec && j >= JscB && j <= g.jec);
  if (!fx && !fy) return;
  const double cxq = fx ? cx[q] : 0.0, cyq = fy ? cy[q] : 0.0;
  for (int m = 0; m < tt.n; m++) {
    double *dx = df.dfx[m], *dy = df.dfy[m];
    if (!dx && !dy) continue;
    const double *t = tt.src[m];
    const double c = t[q];
    if (dx && fx) dx[q] = dx[q] + cxq * (c - t[q + 1]) * Idt;
    if (dy && fy) dy[q] = dy[q] + cyq * (c - t[q + g.P]) * Idt;
  }
}

// df2d_x / df2d_y (:584-591): the same terms summed over k into a 2-D field, k ascending as the reference's loop nest
// visits them (one thread walks the column of a face).
__global__ void __launch_bounds__(HD_TX * HD_TY)
k_hd_df2(Geom g, HdParams p, const __grid_constant__ TrTable tt, const __grid_constant__ DfTable df,
         const double *__restrict__ cx, const double *__restrict__ cy, double Idt, int IscB, int JscB) {
  const int col = blockIdx.x * HD_TX + threadIdx.x, row = blockIdx.y * HD_TY + threadIdx.y;
  if (col >= g.P || row >= g.R) return;
  const int i = col - g.coff, j = row - g.roff;
  const long long q2 = (long long)row * g.P + col;
  const bool fx = (j >= g.jsc && j <= g.jec && i >= IscB && i <= g.iec);
  const bool fy = (i >= g.isc && i <= g.iec && j >= JscB && j <= g.jec);
  if (!fx && !fy) return;
  for (int m = 0; m < tt.n; m++) {
    double *dx = df.d2x[m], *dy = df.d2y[m];
    if (!dx && !dy) continue;
    const double *t = tt.src[m];
    double ax = (dx && fx) ? dx[q2] : 0.0, ay = (dy && fy) ? dy[q2] : 0.0;
    for (int k = 0; k < g.nk; k++) {
      double scale;
      if (!hd_level(p, k + 1, scale)) continue;
      const long long q = q2 + g.plane * k;
      const double c = t[q];
      if (dx && fx) ax = ax + cx[q] * (c - t[q + 1]) * Idt;
      if (dy && fy) ay = ay + cy[q] * (c - t[q + g.P]) * Idt;
    }
    if (dx && fx) dx[q2] = ax;
    if (dy && fy) dy[q2] = ay;
  }
}

void launch_hd_df(cudaStream_t s, const Geom &g, const HdParams &p, const TrTable &tt, const DfTable &df, bool any3,
                  bool any2, const double *cx, const double *cy, double Idt) {
  const int IscB = (g.IsdB == g.isd - 1) ? g.isc - 1 : g.isc, JscB = (g.JsdB == g.jsd - 1) ? g.jsc - 1 : g.jsc;
  dim3 block(HD_TX, HD_TY);
  if (any3) {
    dim3 grid((g.P + HD_TX - 1) / HD_TX, (g.R + HD_TY - 1) / HD_TY, g.nk);
    k_hd_df3<<<grid, block, 0, s>>>(g, p, tt, df, cx, cy, Idt, IscB, JscB);
  }
  if (any2) {
    dim3 grid((g.P + HD_TX - 1) / HD_TX, (g.R + HD_TY - 1) / HD_TY, 1);
    k_hd_df2<<<grid, block, 0, s>>>(g, p, tt, df, cx, cy, Idt, IscB, JscB);
  }
}

#define HD_W 64
#define HD_H 8
#define HD_NS 6

void launch_hd_coef(cudaStream_t s, const Geom &g, const HdParams &p, const double *h, const double *khdt_x,
                    const double *khdt_y, const double *IareaT, double *cx, double *cy, double *ihd) {
  dim3 grid((g.P + HD_TX - 1) / HD_TX, (g.R + HD_TY - 1) / HD_TY, g.nk), block(HD_TX, HD_TY);
  k_hd_coef<<<grid, block, 0, s>>>(g, p, h, khdt_x, khdt_y, IareaT, cx, cy, ihd);
}

void launch_hd_apply(cudaStream_t s, const Geom &g, const HdParams &p, const TrTable &tt, const double *cx,
                     const double *cy, const double *ihd) {
  const int ncol = g.ied - g.isd + 1 + 1;       // + the possible odd leading column (tiles start on an even column)
  const size_t smem = sizeof(double) * HD_NS * (HD_H + 2) * (HD_W + 4);      // 32 640 B: under the 48 KB default
  dim3 grid((ncol + HD_W - 1) / HD_W, (g.R + HD_H - 1) / HD_H, g.nk), block(HD_W, HD_H);
  k_hd_tile<HD_W, HD_H, HD_NS><<<grid, block, smem, s>>>(g, p, tt, cx, cy, ihd);
}
