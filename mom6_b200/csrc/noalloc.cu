// mom6adv_zint100_noalloc: the 100 m column integral of post_tracer_transport_diagnostics
// (/root/reference/src/tracer/MOM_tracer_registry.F90:745-766,805-811; same arithmetic as k_col_khi / k_col_zint100 of
// step.cu) for callers that may not allocate: code under CUDA graph capture, tight serving loops, harnesses that
// forbid cudaMalloc.  This entry point has NO counterpart in the reference.
// It creates no plan and calls no cudaMalloc / cudaFree / allocating library routine: inputs and outputs
// are the caller's device arrays, and the scratch the reduction needs (the per-CTA maxima of "last layer with weight
// under 100 m" and their global maximum khi) lives in the file-scope device array below.
//
//   * The array is a fixed 256 KiB + 256 B that every GPU which loads the library holds until exit.
//   * Limit: ni*nj <= MOM6ADV_NOALLOC_MAX_COLUMNS (8 388 608 columns; the 1/8 degree global grid has 6.2 M).
//   * In host code `g_na_scratch` is a host-side stand-in, not a device address: the address a kernel may use comes
//     from cudaGetSymbolAddress, resolved here once per device with that device current, return code checked.
//   * Every call shares the array and it keeps the last call's contents.  Nothing in it has to start at zero: each
//     call overwrites every slot it later reads (per-CTA maxima [1, 1+nblocks), then slot 0).  Calls that OVERLAP on the
//     GPU (different streams, or graph branches without a dependency) race on it: serialise them.
//   * The call is asynchronous on `stream`; the first call on a device resolves the symbol address (do it once before
//     starting a stream capture).
#include <cuda_runtime.h>
#include <cstdio>
#include <mutex>

#include "../../include/mom6adv.h"

#define NA_TPB 128
#define NA_MAX_BLOCKS 65536
static_assert((long long)NA_TPB * NA_MAX_BLOCKS == MOM6ADV_NOALLOC_MAX_COLUMNS, "header limit out of step with the scratch array");

// slot 0: khi; slots 64.. : one maximum per CTA of the first phase (slot 0 has its own 256-byte line)
static __device__ __align__(256) int g_na_scratch[64 + NA_MAX_BLOCKS];

// weight of layer k under 100 m from the running interface positions (MOM_tracer_registry.F90:750-760)
__device__ __forceinline__ double na_frac100(double ztop, double zbot) {
  if (zbot <= 100.0) return 1.0;
  if (ztop < 100.0) return (100.0 - ztop) / (zbot - ztop);
  return 0.0;
}

// phase 1: per column the last layer with a non-zero weight (:745-766), maximum per CTA -> scratch[64 + block]
__global__ void __launch_bounds__(NA_TPB)
k_na_khi(const double *__restrict__ h, long long ncol, int nk, double H_to_m, int *__restrict__ scratch) {
  const long long c = (long long)blockIdx.x * NA_TPB + threadIdx.x;
  int last = 0;
  if (c < ncol) {
    double zbot = 0.0;
    for (int k = 0; k < nk; k++) {
      const double ztop = zbot;
      zbot = ztop + h[c + ncol * k] * H_to_m;
      if (na_frac100(ztop, zbot) > 0.0) last = k + 1;
    }
  }
  for (int o = 16; o > 0; o >>= 1) last = max(last, __shfl_xor_sync(0xffffffffu, last, o));
  __shared__ int ws[NA_TPB / 32];
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = last;
  __syncthreads();
  if (threadIdx.x == 0) {
    int m = ws[0];
    for (int w = 1; w < NA_TPB / 32; w++) m = max(m, ws[w]);
    scratch[64 + blockIdx.x] = m;
  }
}

// phase 2: one CTA folds the per-CTA maxima into scratch[0] (the reference's khi)
__global__ void __launch_bounds__(1024)
k_na_fold(int nblocks, int *__restrict__ scratch) {
  int m = 0;
  for (int b = threadIdx.x; b < nblocks; b += blockDim.x) m = max(m, scratch[64 + b]);
  for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
  __shared__ int ws[32];
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    int r = ws[0];
    for (int w = 1; w < (int)(blockDim.x >> 5); w++) r = max(r, ws[w]);
    scratch[0] = r;
  }
}

// phase 3: weights recomputed on the fly, sum over k <= khi in the reference's order (:805-811)
__global__ void __launch_bounds__(NA_TPB)
k_na_zint100(const double *__restrict__ h, const double *__restrict__ t, long long ncol, double H_to_m,
             const int *__restrict__ scratch, double *__restrict__ out) {
  const long long c = (long long)blockIdx.x * NA_TPB + threadIdx.x;
  if (c >= ncol) return;
  const int kmax = scratch[0];
  double w = 0.0, zbot = 0.0;
  for (int k = 0; k < kmax; k++) {
    const long long q = c + ncol * k;
    const double ztop = zbot;
    zbot = ztop + h[q] * H_to_m;
    w = w + na_frac100(ztop, zbot) * ((h[q] * H_to_m) * t[q]);
  }
  out[c] = w;
}

#define NA_MAX_DEVICES 64
static int *g_na_ptr[NA_MAX_DEVICES];      // device address of g_na_scratch on each device (nullptr: not resolved yet)
static std::mutex g_na_mutex;

static int na_fail(char *errbuf, int errlen, int code, const char *msg) {
  if (errbuf && errlen > 0) snprintf(errbuf, (size_t)errlen, "%s", msg);
  return code;
}

static bool na_is_device(const void *p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

extern "C" int mom6adv_zint100_noalloc(const double *h_diag, const double *t, int ni, int nj, int nk, double H_to_m,
                                       double *out2d, void *stream, char *errbuf, int errbuf_len) {
  if (errbuf && errbuf_len > 0) errbuf[0] = 0;
  if (!h_diag || !t || !out2d) return na_fail(errbuf, errbuf_len, MOM6ADV_ERR_INVALID, "zint100_noalloc: NULL array");
  if (ni <= 0 || nj <= 0 || nk <= 0) return na_fail(errbuf, errbuf_len, MOM6ADV_ERR_INVALID, "zint100_noalloc: ni, nj, nk must be > 0");
  const long long ncol = (long long)ni * nj;
  if (ncol > MOM6ADV_NOALLOC_MAX_COLUMNS) {
    char b[200];
    snprintf(b, sizeof b, "zint100_noalloc: ni*nj = %lld columns exceed the fixed scratch of %lld columns "
             "(MOM6ADV_NOALLOC_MAX_COLUMNS); use mom6adv_column_diag", ncol, (long long)MOM6ADV_NOALLOC_MAX_COLUMNS);
    return na_fail(errbuf, errbuf_len, MOM6ADV_ERR_UNSUPPORTED, b);
  }
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return na_fail(errbuf, errbuf_len, MOM6ADV_ERR_CUDA, "zint100_noalloc: no CUDA device (there is no CPU fallback)"); }
  if (dev < 0 || dev >= NA_MAX_DEVICES) return na_fail(errbuf, errbuf_len, MOM6ADV_ERR_UNSUPPORTED, "zint100_noalloc: device ordinal out of range");
  if (!na_is_device(h_diag) || !na_is_device(t) || !na_is_device(out2d))
    return na_fail(errbuf, errbuf_len, MOM6ADV_ERR_INVALID, "zint100_noalloc: h_diag, t and out2d must be device arrays");
  int *scratch = nullptr;
  {
    // once per device, with that device current: the array's NAME is only a host-side handle
    std::lock_guard<std::mutex> lk(g_na_mutex);
    if (!g_na_ptr[dev]) {
      void *p = nullptr;
      const cudaError_t e = cudaGetSymbolAddress(&p, g_na_scratch);
      if (e != cudaSuccess || !p) {
        char b[200];
        snprintf(b, sizeof b, "zint100_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
        cudaGetLastError();
        return na_fail(errbuf, errbuf_len, MOM6ADV_ERR_CUDA, b);
      }
      g_na_ptr[dev] = (int *)p;
    }
    scratch = g_na_ptr[dev];
  }
  cudaStream_t s = (cudaStream_t)stream;
  const int nblocks = (int)((ncol + NA_TPB - 1) / NA_TPB);
  k_na_khi<<<nblocks, NA_TPB, 0, s>>>(h_diag, ncol, nk, H_to_m, scratch);
  k_na_fold<<<1, 1024, 0, s>>>(nblocks, scratch);
  k_na_zint100<<<nblocks, NA_TPB, 0, s>>>(h_diag, t, ncol, H_to_m, scratch, out2d);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    char b[200];
    snprintf(b, sizeof b, "zint100_noalloc: kernel launch failed: %s", cudaGetErrorString(e));
    return na_fail(errbuf, errbuf_len, MOM6ADV_ERR_CUDA, b);
  }
  return MOM6ADV_OK;
}
