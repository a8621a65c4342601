// Microbenchmark that settles the B200 (sm_100a) FP64 issue model:
//   does an FP64 warp instruction hold its scheduler's ISSUE PORT for two cycles (cost = nonFP64 + 2*FP64),
//   or only the 16-lane FP64 PIPE (cost = max(all instructions, 2*FP64))?
// Independent DADD/DMUL chains (no FMA, no memory) are interleaved with independent filler instructions of
// another pipe at ratios R = 0, 1, 2, 3 fillers per FP64 instruction, at 1..16 warps per scheduler.
// Time is read with clock64 (SM cycles) around the loop, max over the warps of one CTA.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o issue_model issue_model.cu
//   ./issue_model            -> table on stdout (cycles per FP64 warp instruction per scheduler)
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CHAINS 8

enum { F_NONE = 0, F_IADD, F_LOP, F_FSEL, F_FADD, F_IMAD, F_SHFL, F_LDS, F_DSETP };

template <int KIND>
__device__ __forceinline__ void filler(const int ALT, unsigned &x, unsigned y, float &f, unsigned saddr) {
  // IADD / LOP3 alternate on a chain so that ptxas cannot fold two of them into one three-input instruction
  if (KIND == F_IADD && ALT == 0) asm volatile("add.u32 %0, %0, %1;" : "+r"(x) : "r"(y));
  else if (KIND == F_IADD) asm volatile("lop3.b32 %0, %0, %1, %1, 0x1e;" : "+r"(x) : "r"(y));
  else if (KIND == F_LOP && ALT == 0) asm volatile("lop3.b32 %0, %0, %1, %1, 0x1e;" : "+r"(x) : "r"(y));
  else if (KIND == F_LOP) asm volatile("shf.l.wrap.b32 %0, %0, %0, 3;" : "+r"(x));
  else if (KIND == F_FSEL) asm volatile("{.reg .pred p; setp.ne.u32 p, %2, 0; selp.f32 %0, %0, %1, p;}" : "+f"(f) : "f"(1.5f), "r"(y));
  else if (KIND == F_FADD) asm volatile("add.rn.f32 %0, %0, %1;" : "+f"(f) : "f"(1.25f));
  else if (KIND == F_IMAD) asm volatile("mad.lo.u32 %0, %0, %1, %1;" : "+r"(x) : "r"(y));
  else if (KIND == F_SHFL) asm volatile("shfl.sync.down.b32 %0, %0, 1, 0x1f, 0xffffffff;" : "+r"(x));
  else if (KIND == F_LDS) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x) : "r"(saddr + (x & 0x7c)));
}

// FPK: 0 = DADD/DMUL alternating, 1 = DFMA, 2 = DADD only
template <int KIND, int R, int FPK>
__global__ void k_issue(double *out, long long *cyc, int iters, double c0, double c1) {
  __shared__ unsigned sm[64];
  if (threadIdx.x < 64) sm[threadIdx.x] = threadIdx.x * 4;
  __syncthreads();
  unsigned saddr = (unsigned)__cvta_generic_to_shared(sm);
  double d[CHAINS];
  unsigned x[CHAINS];
  float f[CHAINS];
#pragma unroll
  for (int c = 0; c < CHAINS; c++) { d[c] = 1.0 + 1e-3 * (threadIdx.x + c); x[c] = threadIdx.x * 7 + c; f[c] = 1.0f + c; }
  const unsigned y = (unsigned)iters | 1u;
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int rep = 0; rep < 2; rep++) {
#pragma unroll
      for (int c = 0; c < CHAINS; c++) {
        if (FPK == 1) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d[c]) : "d"(c1), "d"(c0));
        else if (FPK == 2 || rep == 0) asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(d[c]) : "d"(c0));
        else asm volatile("mul.rn.f64 %0, %0, %1;" : "+d"(d[c]) : "d"(c1));
#pragma unroll
        for (int r = 0; r < R; r++) filler<KIND>((r + rep) & 1, x[(c + r) % CHAINS], y, f[(c + r) % CHAINS], saddr);
      }
    }
  }
  const long long t1 = clock64();
  double s = 0; unsigned xs = 0; float fs = 0;
#pragma unroll
  for (int c = 0; c < CHAINS; c++) { s += d[c]; xs ^= x[c]; fs += f[c]; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + xs + fs;
  // max over the CTA's warps of (end - start): one atomicMax per warp
  if ((threadIdx.x & 31) == 0) atomicMax((unsigned long long *)&cyc[blockIdx.x], (unsigned long long)(t1 - t0));
}

// dependent-chain latency: one warp, one chain
template <int FPK>
__global__ void k_latency(double *out, long long *cyc, int iters, double c0, double c1) {
  double d = 1.0 + threadIdx.x;
  const long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int u = 0; u < 16; u++) {
      if (FPK == 1) asm volatile("fma.rn.f64 %0, %0, %1, %2;" : "+d"(d) : "d"(c1), "d"(c0));
      else if (FPK == 2) asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(d) : "d"(c0));
      else asm volatile("mul.rn.f64 %0, %0, %1;" : "+d"(d) : "d"(c1));
    }
  }
  const long long t1 = clock64();
  out[threadIdx.x] = d;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}

static double *d_out;
static long long *d_cyc;
static int n_sm;

template <int KIND, int R, int FPK>
static void run(const char *name, int warps_per_sched) {
  // warps_per_sched warps on each of the 4 schedulers of every SM: CTAs of up to 512 threads (one per 4 warps/scheduler)
  const int wps_cta = warps_per_sched < 4 ? warps_per_sched : 4;
  const int ctas_per_sm = warps_per_sched / wps_cta;
  const int threads = 32 * 4 * wps_cta;
  const int grid = n_sm * ctas_per_sm;
  const int iters = 4000;
  cudaMemset(d_cyc, 0, sizeof(long long) * grid);
  k_issue<KIND, R, FPK><<<grid, threads>>>(d_out, d_cyc, 200, 1e-9, 1.0000001);      // warm-up
  cudaMemset(d_cyc, 0, sizeof(long long) * grid);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k_issue<KIND, R, FPK><<<grid, threads>>>(d_out, d_cyc, iters, 1e-9, 1.0000001);
  cudaEventRecord(e1);
  cudaError_t err = cudaDeviceSynchronize();
  if (err != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(err)); exit(1); }
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long *h = (long long *)malloc(sizeof(long long) * grid);
  cudaMemcpy(h, d_cyc, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
  double mean = 0; long long mx = 0;
  for (int i = 0; i < grid; i++) { mean += h[i]; if (h[i] > mx) mx = h[i]; }
  mean /= grid;
  free(h);
  // FP64 warp instructions per scheduler in the loop: warps_per_sched * iters * 2*CHAINS
  const double nf = (double)warps_per_sched * iters * 2 * CHAINS;
  printf("%-6s fp=%d R=%d w/sched=%2d  cycles/FP64instr/sched: mean %.3f  max %.3f   (sum model %d, max model %d)  %.3f ms\n",
         name, FPK, R, warps_per_sched, mean / nf, (double)mx / nf, 2 + R, (1 + R) > 2 ? 1 + R : 2, ms);
}

template <int KIND, int FPK>
static void sweep(const char *name) {
  const int ws[] = {1, 2, 4, 8, 16};
  for (int w : ws) run<KIND, 0, FPK>(name, w);
  for (int w : ws) run<KIND, 1, FPK>(name, w);
  for (int w : ws) run<KIND, 2, FPK>(name, w);
  for (int w : ws) run<KIND, 3, FPK>(name, w);
}

int main(int argc, char **argv) {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  n_sm = p.multiProcessorCount;
  printf("device %s, %d SMs, cc %d.%d\n", p.name, n_sm, p.major, p.minor);
  cudaMalloc(&d_out, sizeof(double) * 2048 * n_sm * 4);
  cudaMalloc(&d_cyc, sizeof(long long) * n_sm * 8);
  if (argc > 1) {      // short list for an ncu counter pass (smsp__issue_active, sm__pipe_fp64_cycles_active)
    run<F_FSEL, 0, 0>("FSEL", 8); run<F_FSEL, 1, 0>("FSEL", 8); run<F_FSEL, 2, 0>("FSEL", 8); run<F_FSEL, 3, 0>("FSEL", 8);
    run<F_IMAD, 1, 0>("IMAD", 8); run<F_SHFL, 1, 0>("SHFL", 8);
    return 0;
  }
  // dependent latencies
  {
    const int iters = 2000; long long c;
    k_latency<0><<<1, 32>>>(d_out, d_cyc, iters, 1e-9, 1.0000001); cudaMemcpy(&c, d_cyc, 8, cudaMemcpyDeviceToHost);
    printf("latency DMUL dependent: %.2f cycles\n", (double)c / (iters * 16));
    k_latency<2><<<1, 32>>>(d_out, d_cyc, iters, 1e-9, 1.0000001); cudaMemcpy(&c, d_cyc, 8, cudaMemcpyDeviceToHost);
    printf("latency DADD dependent: %.2f cycles\n", (double)c / (iters * 16));
    k_latency<1><<<1, 32>>>(d_out, d_cyc, iters, 1e-9, 1.0000001); cudaMemcpy(&c, d_cyc, 8, cudaMemcpyDeviceToHost);
    printf("latency DFMA dependent: %.2f cycles\n", (double)c / (iters * 16));
  }
  sweep<F_IADD, 0>("IADD");
  sweep<F_LOP, 0>("LOP3");
  sweep<F_FSEL, 0>("FSEL");
  sweep<F_FADD, 0>("FADD");
  sweep<F_IMAD, 0>("IMAD");
  sweep<F_SHFL, 0>("SHFL");
  sweep<F_LDS, 0>("LDS");
  // DFMA for reference (the 37 TFLOP/s pipe rate)
  { const int ws[] = {1, 4, 8, 16}; for (int w : ws) run<F_IADD, 0, 1>("IADD", w); for (int w : ws) run<F_IADD, 1, 1>("IADD", w); }
  return 0;
}
