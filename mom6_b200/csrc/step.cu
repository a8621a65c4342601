// The two small loops either side of advect_tracer that decide whether its inputs and outputs have to cross PCIe every
// step: the accumulation of the tracer transports during the dynamics steps and the
// column reductions of post_tracer_transport_diagnostics.  With them on the device, uhtr / vhtr never exist on the host
// and a diagnostic costs one 2-D field of D2H instead of a 3-D one.
//
//   k_accumulate   /root/reference/src/core/MOM_dynamics_split_RK2.F90:1045-1052   uhtr += uh*dt, vhtr += vh*dt
//   (zeroing        /root/reference/src/core/MOM.F90:1465-1466 is a memset)
//   k_col_sum      /root/reference/src/tracer/MOM_tracer_registry.F90:786-791       sum_k advection_xy
//   k_col_zint     :796-802                                                         sum_k (h*H_to_m)*t
//   k_col_khi      :745-766   last layer with a non-zero weight under 100 m anywhere on this PE
//   k_col_zint100  :745-764,805-811   weights recomputed on the fly, sum over k <= khi
// All sums run over k in the reference's order inside one thread per column: bitwise the Fortran loop.
#include "internal.cuh"
#include "kernels.h"

__global__ void k_accumulate(Geom g, double *__restrict__ acc, const double *__restrict__ add, double dt,
                             int ilo, int ihi, int jlo, int jhi) {
  const int ci = blockIdx.x * blockDim.x + threadIdx.x, rj = blockIdx.y, k0 = blockIdx.z;
  if (ci >= g.P) return;
  const int i = ci - g.coff, j = rj - g.roff;
  if (i < ilo || i > ihi || j < jlo || j > jhi) return;
  const long long q = (long long)ci + (long long)g.P * rj + g.plane * k0;
  acc[q] = acc[q] + add[q] * dt;
}

__device__ __forceinline__ bool col_of(const Geom &g, int &ci, int &rj) {
  ci = blockIdx.x * blockDim.x + threadIdx.x; rj = blockIdx.y;
  if (ci >= g.P) return false;
  const int i = ci - g.coff, j = rj - g.roff;
  return i >= g.isc && i <= g.iec && j >= g.jsc && j <= g.jec;
}

__global__ void k_col_sum(Geom g, const double *__restrict__ a, double *__restrict__ out) {
  int ci, rj;
  const bool in = col_of(g, ci, rj);
  if (ci >= g.P) return;
  const long long q2 = (long long)ci + (long long)g.P * rj;
  double w = 0.0;
  if (in) for (int k = 0; k < g.nk; k++) w = w + a[q2 + g.plane * k];
  out[q2] = w;
}

__global__ void k_col_zint(Geom g, const double *__restrict__ h, const double *__restrict__ t, double H_to_m,
                           double *__restrict__ out) {
  int ci, rj;
  const bool in = col_of(g, ci, rj);
  if (ci >= g.P) return;
  const long long q2 = (long long)ci + (long long)g.P * rj;
  double w = 0.0;
  if (in) for (int k = 0; k < g.nk; k++) { const long long q = q2 + g.plane * k; w = w + (h[q] * H_to_m) * t[q]; }
  out[q2] = w;
}

// weight of layer k under 100 m from the running interface positions (:750-760)
__device__ __forceinline__ double frac100(double ztop, double zbot) {
  if (zbot <= 100.0) return 1.0;
  if (ztop < 100.0) return (100.0 - ztop) / (zbot - ztop);
  return 0.0;
}

__global__ void k_col_khi(Geom g, const double *__restrict__ h, double H_to_m, int *__restrict__ khi) {
  int ci, rj;
  if (!col_of(g, ci, rj)) return;
  const long long q2 = (long long)ci + (long long)g.P * rj;
  double zbot = 0.0;
  int last = 0;
  for (int k = 0; k < g.nk; k++) {
    const double ztop = zbot;
    zbot = ztop + h[q2 + g.plane * k] * H_to_m;
    if (frac100(ztop, zbot) > 0.0) last = k + 1;
  }
  if (last > 0) atomicMax(khi, last);
}

__global__ void k_col_zint100(Geom g, const double *__restrict__ h, const double *__restrict__ t, double H_to_m,
                              const int *__restrict__ khi, double *__restrict__ out) {
  int ci, rj;
  const bool in = col_of(g, ci, rj);
  if (ci >= g.P) return;
  const long long q2 = (long long)ci + (long long)g.P * rj;
  double w = 0.0;
  if (in) {
    const int kmax = *khi;
    double zbot = 0.0;
    for (int k = 0; k < kmax; k++) {
      const long long q = q2 + g.plane * k;
      const double ztop = zbot;
      zbot = ztop + h[q] * H_to_m;
      w = w + frac100(ztop, zbot) * ((h[q] * H_to_m) * t[q]);
    }
  }
  out[q2] = w;
}

void launch_accumulate(cudaStream_t s, const Geom &g, double *acc, const double *add, double dt, int ilo, int ihi,
                       int jlo, int jhi) {
  dim3 grid((g.P + 127) / 128, g.R, g.nk);
  k_accumulate<<<grid, 128, 0, s>>>(g, acc, add, dt, ilo, ihi, jlo, jhi);
}
void launch_col_sum(cudaStream_t s, const Geom &g, const double *a, double *out) {
  k_col_sum<<<dim3((g.P + 127) / 128, g.R, 1), 128, 0, s>>>(g, a, out);
}
void launch_col_zint(cudaStream_t s, const Geom &g, const double *h, const double *t, double H_to_m, double *out) {
  k_col_zint<<<dim3((g.P + 127) / 128, g.R, 1), 128, 0, s>>>(g, h, t, H_to_m, out);
}
void launch_col_zint100(cudaStream_t s, const Geom &g, const double *h, const double *t, double H_to_m, int *khi,
                        double *out) {
  cudaMemsetAsync(khi, 0, sizeof(int), s);
  k_col_khi<<<dim3((g.P + 127) / 128, g.R, 1), 128, 0, s>>>(g, h, H_to_m, khi);
  k_col_zint100<<<dim3((g.P + 127) / 128, g.R, 1), 128, 0, s>>>(g, h, t, H_to_m, khi, out);
}
