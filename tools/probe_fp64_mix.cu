// tools/probe_fp64_mix.cu -- what caps a DFMA-heavy kernel (K1) below the DFMA issue peak?
//   mode 0: DFMA, two register sources shared by all chains (the tools/probe_fp64.cu figure)
//   mode 1: DFMA, one source from the constant bank (c[0x3][...] operand)
//   mode 2: DFMA with three distinct register operands per instruction (48 live doubles)
//   mode 3: mode 0 + one independent integer instruction per DFMA
//   mode 4: mode 0 + one independent integer instruction per two DFMA
//   mode 5: mode 2 + one integer instruction per two DFMA
// WARPS per CTA is a template parameter (K1 runs 4 warps per scheduler).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o probe_fp64_mix tools/probe_fp64_mix.cu
#include <cstdio>
#include <cuda_runtime.h>

__constant__ double kc[16];

template <int MODE>
__global__ void __launch_bounds__(256) probe(double* out, int iters, double a0, double b0, int i0) {
  double f[16], g[16], h[16];
  int n[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    f[i] = threadIdx.x * 1e-9 + i;
    g[i] = 1.0 + (threadIdx.x + i) * 1e-12;
    h[i] = 1e-9 * (i + 1);
    n[i] = i0 + i;
  }
  const double a = a0 + threadIdx.x * 1e-12, b = b0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int rep = 0; rep < 2; ++rep) {
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        if (MODE == 0 || MODE == 3 || MODE == 4) asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(f[i]) : "d"(a), "d"(b));
        if (MODE == 1) f[i] = fma(f[i], kc[i], b);
        if (MODE == 2 || MODE == 5) asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(f[i]) : "d"(g[i]), "d"(h[(i + rep) & 15]));
        if (MODE == 3 || ((MODE == 4 || MODE == 5) && (i & 1)))
          asm volatile("add.s32 %0, %0, %1;" : "+r"(n[i]) : "r"(i0));
      }
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += f[i] + g[i] + h[i] + n[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
static void run(const char* name, int ctas_per_sm, int threads) {
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int grid = sms * ctas_per_sm, iters = 20000;
  double* out; cudaMalloc(&out, (size_t)grid * 256 * 8);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  probe<MODE><<<grid, threads>>>(out, 200, 1.0, 1e-9, 1);
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < 3; ++r) {
    cudaEventRecord(e0);
    probe<MODE><<<grid, threads>>>(out, iters, 1.0, 1e-9, 1);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  const double flop = (double)grid * threads * iters * 32 * 2.0;
  printf("{\"probe\": \"%s\", \"ctas_per_sm\": %d, \"warps_per_sm\": %d, \"ms\": %.3f, \"dfma_tflops\": %.2f}\n", name,
         ctas_per_sm, ctas_per_sm * threads / 32, best, flop / best / 1e9);
  cudaFree(out);
}

int main() {
  double hk[16];
  for (int i = 0; i < 16; ++i) hk[i] = 1.0 - 1e-9 * i;
  cudaMemcpyToSymbol(kc, hk, sizeof(hk));
  for (int c : {2, 4}) {
    run<0>("dfma reg,reg", c, 128);
    run<1>("dfma reg,const", c, 128);
    run<2>("dfma 3 distinct regs", c, 128);
    run<3>("dfma + 1 int per dfma", c, 128);
    run<4>("dfma + 1 int per 2 dfma", c, 128);
    run<5>("dfma 3 distinct regs + 1 int per 2 dfma", c, 128);
  }
  return 0;
}
