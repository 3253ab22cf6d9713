// tools/probe_umma.cu -- what bounds the int8 tcgen05 contraction?  Issue rate of kind::i8 (and, for comparison,
// kind::f8f6f4) MMAs of the shapes the slice contraction uses, operands resident in shared memory (no loads), for the
// no-swizzle and the 128-byte-swizzle K-major layouts.  One CTA per SM, one issuing thread, completion by
// tcgen05.commit on two alternating mbarriers (the issuer runs one group ahead).  Data is irrelevant (timing only).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/_bin/probe_umma tools/probe_umma.cu
// Output: one JSON object per line (-> profiles/r02_probe_umma.jsonl).
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "../gp_dla_detection_b200/csrc/slices.cuh"

using namespace dla::sl;

#define CK(x)                                                                                   \
  do {                                                                                          \
    cudaError_t e_ = (x);                                                                       \
    if (e_ != cudaSuccess) {                                                                    \
      fprintf(stderr, "CUDA error %s at %s:%d: %s\n", #x, __FILE__, __LINE__, cudaGetErrorString(e_)); \
      exit(2);                                                                                  \
    }                                                                                           \
  } while (0)

__device__ __forceinline__ void umma_f8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// A operand from tensor memory: one 128 x 256 bit copy (128 rows x 32 bytes = one slice of a K step) from shared memory,
// then MMAs that name the TMEM address instead of a shared-memory descriptor
__device__ __forceinline__ void tmem_cp_128x256b(uint32_t taddr, uint64_t sdesc) {
  asm volatile("tcgen05.cp.cta_group::1.128x256b [%0], %1;" ::"r"(taddr), "l"(sdesc) : "memory");
}
__device__ __forceinline__ void umma_i8_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// A group of MMAs, all operands compile-time offsets from one warp-uniform base (the product kernel's form: the
// descriptors live in uniform registers and the issue loop is UTCIMMA after UTCIMMA).
//   PAT 0: 8 x (128 x N x 32), distinct accumulators     PAT 1: the slice K step of an N-column unit (10 or 7 wide MMAs)
//   PAT 2: the slice K step as 28 separate 128 x N x 32 MMAs
//   PAT 3: PAT 1 with the sample slices copied to tensor memory first (tcgen05.cp) and read from there by the MMAs
//   PAT 4: PAT 3 without the copies (the A operand already in tensor memory): the MMAs alone
template <int PAT, int N, int KIND, int LAYOUT>
__device__ __forceinline__ void issue_group(uint32_t base, uint32_t tmem_base) {
  constexpr uint32_t SBO = LAYOUT == 0 ? 128u : 1024u;
  constexpr uint64_t HI = (uint64_t)(((SBO >> 4) & 0x3FFFu) | (1u << 14) | ((uint32_t)LAYOUT << 29)) << 32;
  constexpr uint32_t LBO_A = LAYOUT == 0 ? 128u * 16u : 16u;
  constexpr uint32_t LBO_B = LAYOUT == 0 ? (PAT == 1 ? 7u * N * 16u : N * 16u) : 16u;
  const uint32_t a_lo = ((base >> 4) & 0x3FFFu) | ((LBO_A >> 4) << 16);
  const uint32_t b_lo = (((base + 131072u) >> 4) & 0x3FFFu) | ((LBO_B >> 4) << 16);
  auto mma = [&](uint32_t dcol, uint32_t aoff, uint32_t boff, int n) {
    const uint64_t ad = HI | (uint64_t)(a_lo + (aoff >> 4)), bd = HI | (uint64_t)(b_lo + (boff >> 4));
    if (KIND == 0) umma_i8(tmem_base + dcol, ad, bd, idesc_i8(128, n), 1u);
    else umma_f8(tmem_base + dcol, ad, bd, (1u << 4) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24), 1u);
  };
  if (PAT == 0) {
#pragma unroll
    for (int j = 0; j < 8; ++j) mma((uint32_t)((j * N) % 512), (uint32_t)(j % 4) * (LAYOUT == 0 ? 4096u : 32u), (uint32_t)(j % 2) * (LAYOUT == 0 ? 8192u : 64u), N);
  } else if (PAT == 1) {
#pragma unroll
    for (int i = 0; i < 7; ++i) {
      const int ntot = (7 - i) * N;
#pragma unroll
      for (int r0 = 0; r0 < ntot; r0 += 256) {
        const int np = ntot - r0 < 256 ? ntot - r0 : 256;
        mma((uint32_t)(i * N + r0), (uint32_t)i * (LAYOUT == 0 ? 4096u : 16384u), (uint32_t)r0 * (LAYOUT == 0 ? 16u : 128u), np);
      }
    }
  } else if (PAT == 3 || PAT == 4) {
    constexpr uint32_t ACOL = 448;   // accumulators [0, 7 N), sample slices [448, 504): 8 columns of 32 bits per slice
#pragma unroll
    for (int i = 0; i < 7; ++i) {
      if (PAT == 3) tmem_cp_128x256b(tmem_base + ACOL + 8u * i, HI | (uint64_t)(a_lo + ((uint32_t)i * 4096u >> 4)));
      const int ntot = (7 - i) * N;
#pragma unroll
      for (int r0 = 0; r0 < ntot; r0 += 256) {
        const int np = ntot - r0 < 256 ? ntot - r0 : 256;
        umma_i8_ts(tmem_base + (uint32_t)(i * N + r0), tmem_base + ACOL + 8u * i,
                   HI | (uint64_t)(b_lo + (((uint32_t)r0 * 16u) >> 4)), idesc_i8(128, np), 1u);
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < 7; ++i)
#pragma unroll
      for (int j = 0; j + i < 7; ++j) mma((uint32_t)((i + j) * N), (uint32_t)i * 4096u, (uint32_t)j * (2u * N * 16u), N);
  }
}

template <int PAT, int N, int KIND, int LAYOUT>
__global__ void __launch_bounds__(128, 1) probe_kernel(int iters, int depth, long long* cycles) {
  extern __shared__ unsigned char raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bars[8];
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x01010101u * (i & 3);
  if (tid == 0) {
    for (int b = 0; b < 8; ++b) bar_init(&bars[b], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&tmem_slot)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  if (warp == 0) {
    const uint32_t base = s32(smem);
    long long t0 = 0;
    for (int it = 0; it < iters + 8; ++it) {
      if (it == 8) t0 = clock64();
      uint32_t pred;
      asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
      if (pred) {
        issue_group<PAT, N, KIND, LAYOUT>(base, tmem_base);
        umma_commit(&bars[it & 7]);
      }
      __syncwarp();
      const int w = it - (depth - 1);
      if (w >= 0) bar_wait(&bars[w & 7], (w >> 3) & 1);
    }
    for (int w = iters + 8 - (depth - 1); w < iters + 8; ++w) bar_wait(&bars[w & 7], (w >> 3) & 1);
    if (tid == 0) cycles[blockIdx.x] = clock64() - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
}

template <int PAT, int N, int KIND, int LAYOUT>
static void run(const char* name, int nmma, double ideal_clk, int depth = 6, int iters = 2000) {
  long long* d;
  CK(cudaMalloc(&d, 148 * 8));
  CK(cudaFuncSetAttribute(probe_kernel<PAT, N, KIND, LAYOUT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 201 * 1024 + 1024));
  probe_kernel<PAT, N, KIND, LAYOUT><<<148, 128, 201 * 1024 + 1024>>>(iters, depth, d);
  CK(cudaDeviceSynchronize());
  std::vector<long long> h(148);
  CK(cudaMemcpy(h.data(), d, 148 * 8, cudaMemcpyDeviceToHost));
  double mx = 0, sum = 0;
  for (long long v : h) { sum += (double)v; if ((double)v > mx) mx = (double)v; }
  const double per = sum / 148 / iters;
  printf("{\"probe\": \"%s\", \"kind\": \"%s\", \"layout\": \"%s\", \"mmas_per_group\": %d, \"groups_in_flight\": %d, "
         "\"clk_per_group\": %.1f, \"clk_per_group_slowest_sm\": %.1f, \"ideal_clk\": %.0f, \"frac_of_ideal_rate\": %.3f}\n",
         name, KIND == 0 ? "i8" : "f8f6f4", LAYOUT == 0 ? "none" : "swizzle128", nmma, depth, per, mx / iters, ideal_clk, ideal_clk / per);
  fflush(stdout);
  CK(cudaFree(d));
}

template <int KIND, int LAYOUT>
static void shapes() {
  run<0, 32, KIND, LAYOUT>("8 x (128 x 32 x 32)", 8, 8.0 * 16);
  run<0, 64, KIND, LAYOUT>("8 x (128 x 64 x 32)", 8, 8.0 * 32);
  run<0, 128, KIND, LAYOUT>("8 x (128 x 128 x 32)", 8, 8.0 * 64);
  run<0, 256, KIND, LAYOUT>("8 x (128 x 256 x 32)", 8, 8.0 * 128);
}

int main() {
  shapes<0, 0>();
  shapes<0, 2>();
  shapes<1, 0>();
  shapes<1, 2>();
  run<1, 64, 0, 0>("slice K step, 64-column unit, wide MMAs", 10, 896);
  run<1, 64, 0, 0>("slice K step, 64-column unit, wide MMAs", 10, 896, 1);
  run<1, 32, 0, 0>("slice K step, 32-column unit, wide MMAs", 7, 448);
  run<1, 64, 0, 2>("slice K step, 64-column unit, wide MMAs", 10, 896);
  run<1, 32, 0, 2>("slice K step, 32-column unit, wide MMAs", 7, 448);
  run<2, 64, 0, 0>("slice K step, 64-column unit, 28 separate MMAs", 28, 896);
  run<2, 32, 0, 0>("slice K step, 32-column unit, 28 separate MMAs", 28, 448);
  run<1, 64, 1, 0>("slice K step, 64-column unit, wide MMAs", 10, 896);
  run<3, 64, 0, 0>("slice K step, 64-column unit, wide MMAs, A copied to TMEM per slice", 10, 896);
  run<4, 64, 0, 0>("slice K step, 64-column unit, wide MMAs, A resident in TMEM", 10, 896);
  run<3, 32, 0, 0>("slice K step, 32-column unit, wide MMAs, A copied to TMEM per slice", 7, 448);
  run<4, 32, 0, 0>("slice K step, 32-column unit, wide MMAs, A resident in TMEM", 7, 448);
  return 0;
}
