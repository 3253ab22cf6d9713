// tools/probe_fp64.cu -- measures the FP64 denominators MEASURED_PEAKS.json lacks:
//   (a) register-resident DMMA.8x8x4 issue rate  (b) register-resident DFMA issue rate
//   (c) both interleaved in one warp (do they share a pipe?)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o probe_fp64 tools/probe_fp64.cu
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

template <int MODE>  // 0 dmma, 1 dfma, 2 both
__global__ void __launch_bounds__(256) probe(double* out, int iters, double a0, double b0) {
  double c[16][2];
  double f[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) { c[i][0] = c[i][1] = 0.0; f[i] = threadIdx.x * 1e-9 + i; }
  double a = a0 + threadIdx.x * 1e-12, b = b0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      if (MODE == 0 || MODE == 2) dmma(c[i][0], c[i][1], a, b);
      if (MODE == 1 || MODE == 2) {
        asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(f[i]) : "d"(a), "d"(b));
        if (MODE == 1) asm volatile("fma.rn.f64 %0, %1, %2, %0;" : "+d"(f[i]) : "d"(b), "d"(a));
      }
    }
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += c[i][0] + c[i][1] + f[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
static void run(const char* name, int ctas_per_sm, double flop_per_thread_iter) {
  int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int grid = sms * ctas_per_sm, iters = 20000;
  double* out; cudaMalloc(&out, (size_t)grid * 256 * 8);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  probe<MODE><<<grid, 256>>>(out, 200, 1.0, 1e-9);
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < 3; ++r) {
    cudaEventRecord(e0);
    probe<MODE><<<grid, 256>>>(out, iters, 1.0, 1e-9);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
  }
  const double flop = (double)grid * 256 * iters * flop_per_thread_iter;
  printf("{\"probe\": \"%s\", \"ctas_per_sm\": %d, \"ms\": %.3f, \"tflops\": %.2f}\n", name, ctas_per_sm, best,
         flop / best / 1e9);
  cudaFree(out);
}

int main() {
  cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
  printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz\": %d}\n", p.name, p.multiProcessorCount, p.clockRate);
  // per thread per iter: 16 DMMA * 512 flop / 32 threads = 256 flop ; DFMA: 32 fma * 2 = 64 flop
  for (int c : {1, 2, 4}) run<0>("dmma_8x8x4", c, 16 * 512.0 / 32);
  for (int c : {1, 2, 4}) run<1>("dfma", c, 32 * 2.0);
  for (int c : {2}) run<2>("dmma+dfma(16 each)", c, 16 * 512.0 / 32 + 16 * 2.0);
  return 0;
}
