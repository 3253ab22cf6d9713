// slices.cuh -- the Woodbury contraction G = [W | U] . [Phi | M] (log_mvnpdf_low_rank.m:22 and the M'D^-1 y part
// of :26-28) taken OFF the FP64 pipe: an exact-slice ("Ozaki-type") integer contraction on the 5th-generation
// tensor cores.
//
// sm_100a has no f64 kind for tcgen05 (the FP64 tensor path is the warp-level DMMA that shares one pipe with the
// Voigt kernel's DFMA work), but it has kind::i8 at 8192 MAC / clk / SM.  Every fp64 operand is written as NS
// balanced base-256 digits (signed 8-bit "slices"),
//     x = sigma * sum_{i < NS} a_i 2^(-8 i - 6),   a_i in [-128, 127],
// the products of slices i (A side) and j (B side) with i + j = d share one scale, so they accumulate EXACTLY in
// one int32 TMEM accumulator per diagonal d (|acc_d| <= (d+1) K 2^14 < 2^31 for K <= 16 000 pixels); products with
// i + j >= NS are dropped (they are below 2^(-8 NS - 4) of sigma_row tau_col per term).  The epilogue recombines
//     G = tau_col 2^-12 sum_d acc_d 2^(-8 d)            (row scale 1: see below)
// in fp64.  NS = 7 reproduces an fp64-accumulated dot product to about 2^-48 of sum |w||phi|; NS = 8 is better
// than fp64 accumulation.  Measured numbers: DESIGN.md section 6c, profiles/r02_probe_slices.jsonl.
//
// Scaling (no row maximum needed): every pixel p carries a power of two c_p >= the largest value w_sp can take,
// w = a^2 / (omega2 a^2 + nu) <= 1 / (omega2 + nu) for a in [0, 1]; K1 emits w / c_p in [0, 1] and the table side
// holds Phi_p c_p, whose column scale tau_j (a power of two >= max_p |Phi_pj c_p|) is found when the table is built.
//
// Geometry: a CTA tile is SB = 128 samples (UMMA M) x NB = 64 columns x all NS diagonals = NS * 64 <= 512 TMEM
// columns; K step = 32 pixels (UMMA K for 8-bit operands).  Operand planes in HBM are already in the UMMA
// K-major / no-swizzle canonical layout (core matrix = 8 rows x 16 bytes, contiguous), so that one 1-D TMA bulk
// copy per operand brings a whole K step of all NS slices:
//     A planes: [sample block][K step][slice][16-pixel chunk (2)][128 rows][16 B]   = NS * 4 KB per K step
//     B planes: [column block][K step][slice][16-pixel chunk (2)][ 64 rows][16 B]   = NS * 2 KB per K step
// A lane of K1 that owns one sample writes, per 16 pixels and slice, one 16-byte piece; a warp's 32 pieces are
// 512 contiguous bytes.
#ifndef GPDLA_SLICES_CUH
#define GPDLA_SLICES_CUH

#include <stdint.h>
#include <cuda_runtime.h>

namespace dla {
namespace sl {

constexpr int SB = 128;   // samples per tile (UMMA M)
constexpr int KS = 32;    // pixels per K step (UMMA K, 8-bit operands)
constexpr int NB = 64;    // columns per tile (UMMA N)
constexpr int THREADS = 192;  // warp 0 producer, warp 1 MMA issue, warps 2..5 epilogue (TMEM lane quarters 2,3,0,1)

template <int NS>
struct Geo {
  static constexpr int A_STEP = NS * 2 * SB * 16;   // bytes of A planes per K step
  static constexpr int B_STEP = NS * 2 * NB * 16;   // bytes of B planes per K step
  static constexpr int STAGE = A_STEP + B_STEP;
  static constexpr int STAGES = (216 * 1024) / STAGE;   // NS = 6: 6, 7: 5, 8: 4
  static constexpr int TMEM_COLS = 512;
  static constexpr size_t SMEM = (size_t)STAGES * STAGE + 1024 /* alignment slack */ + 256 /* barriers */;
  static_assert(NS >= 2 && NS <= 8, "NS * 64 accumulator columns must fit the 512 TMEM columns");
};

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void bar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void bar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(bar)) : "memory");
}
// bounded: a lost copy / MMA must fault, never hang the GPU
__device__ __forceinline__ void bar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = s32(bar);
  uint32_t done, spins = 0;
  do {
    if (++spins > (1u << 26)) __trap();
    asm volatile(
        "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(dst)),
               "l"(src), "r"(bytes), "r"(s32(bar))
               : "memory");
}

// UMMA shared-memory matrix descriptor, K-major, no swizzle (cute::UMMA::SmemDescriptor):
//   [0,14) start address >> 4   [16,30) leading byte offset >> 4 (between the two 16-byte K chunks)
//   [32,46) stride byte offset >> 4 (between 8-row groups)   [46,48) version = 1   [61,64) layout type 0
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  const uint32_t lo = ((addr >> 4) & 0x3FFFu) | (((lbo >> 4) & 0x3FFFu) << 16);
  const uint32_t hi = ((sbo >> 4) & 0x3FFFu) | (1u << 14);
  return ((uint64_t)hi << 32) | lo;
}
// UMMA instruction descriptor (cute::UMMA::InstrDescriptor), kind::i8: D = S32 (2 at [4,6)), A = B = signed 8 bit
// (1 at [7,10) and [10,13)), both K-major, N >> 3 at [17,23), M >> 4 at [24,29)
__host__ __device__ constexpr uint32_t idesc_i8(int M, int N) {
  return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {  // arrives on bar when every MMA issued so far has completed
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// int32 -> double without the conversion unit: 2^52 + 2^31 + x as a bit pattern, minus the bias (one LOP + one DADD)
__device__ __forceinline__ double i32_to_f64(uint32_t x) {
  return __hiloint2double(0x43300000, (int)(x ^ 0x80000000u)) - 4503601774854144.0;
}

// Balanced base-256 digits of X (|X| <= 2^(8 NS - 1)): byte i (from the top) of the result is slice i as int8.
// X + 0x..808080 (NS - 1 low bytes) carries every digit into [0, 255]; xor 0x80 recentres it to [-128, 127].
template <int NS>
__device__ __forceinline__ uint64_t balanced_digits(long long X) {
  constexpr uint64_t C = (NS >= 8) ? 0x0080808080808080ull : ((0x0080808080808080ull) >> (8 * (8 - NS)));
  return ((uint64_t)X + C) ^ C;
}
// value in [-2, 2] (in units of the scale) -> integer of 8 NS - 2 fractional bits
template <int NS>
__device__ __forceinline__ long long to_fixed(double t) {
  return __double2ll_rn(t * (double)(1ull << (8 * NS - 2)));
}

// ------------------------------------------------------------------------------------------------------------
// The contraction.  Work item t -> sample block sb = t / ncb, column block cb = t % ncb (column blocks of one sample
// block are neighbours in t, so the CTAs that share A run at the same time and its second..fourth read hit L2).
//   Aplanes + sb * a_sb_stride : [K step][slice][2][128][16]
//   Bplanes + cb * b_cb_stride : [K step][slice][2][ 64][16]
//   colscale[cb * 64 + c]      : tau_c 2^-12 (NaN poisons the column, as a NaN table entry must)
//   G[(sb * 128 + r) * ldg + cb * 64 + c]
// ------------------------------------------------------------------------------------------------------------
struct SliceJob {
  const int8_t* Aplanes;
  const int8_t* Bplanes;
  const double* colscale;
  double* G;
  int64_t a_sb_stride, b_cb_stride;
  int nsb, ncb, nks, ldg;
};

template <int NS, int ORDER /* 0: diagonal-major issue order, 1: A-slice-major */>
__global__ void __launch_bounds__(THREADS, 1)
contract_slices_kernel(const SliceJob* __restrict__ jobs, int njobs, const int* __restrict__ tile_ptr /* [njobs+1] */) {
  using G_ = Geo<NS>;
  extern __shared__ unsigned char sl_smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(sl_smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)G_::STAGES * G_::STAGE);
  uint64_t* empty = full + G_::STAGES;
  uint64_t* tfull = empty + G_::STAGES;
  uint64_t* tempty = tfull + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 1);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ntiles = tile_ptr[njobs];

  if (tid == 0) {
    for (int s = 0; s < G_::STAGES; ++s) {
      bar_init(&full[s], 1);
      bar_init(&empty[s], 1);
    }
    bar_init(tfull, 1);
    bar_init(tempty, 4);   // one arrive per epilogue warp
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(tmem_slot)), "n"(G_::TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // work item -> (job, sb, cb)
  auto locate = [&](int t, int& j, int& sb, int& cb) {
    int lo = 0, hi = njobs - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (tile_ptr[mid] <= t) lo = mid; else hi = mid - 1;
    }
    j = lo;
    const int r = t - tile_ptr[lo];
    const int ncb = jobs[lo].ncb;
    sb = r / ncb;
    cb = r - sb * ncb;
  };

  if (warp == 0) {
    // ---------------- producer: one lane streams the K steps of all of this CTA's tiles ----------------
    if (lane == 0) {
      uint32_t it = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        int j, sb, cb;
        locate(t, j, sb, cb);
        const SliceJob job = jobs[j];
        const int8_t* a = job.Aplanes + (int64_t)sb * job.a_sb_stride;
        const int8_t* b = job.Bplanes + (int64_t)cb * job.b_cb_stride;
        for (int ks = 0; ks < job.nks; ++ks, ++it) {
          const int st = it % G_::STAGES;
          bar_wait(&empty[st], ((it / G_::STAGES) & 1) ^ 1);
          unsigned char* dst = smem + (size_t)st * G_::STAGE;
          bar_expect_tx(&full[st], G_::STAGE);
          bulk_load(dst, a + (int64_t)ks * G_::A_STEP, G_::A_STEP, &full[st]);
          bulk_load(dst + G_::A_STEP, b + (int64_t)ks * G_::B_STEP, G_::B_STEP, &full[st]);
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issue: one lane; NS (NS + 1) / 2 integer MMAs per K step ----------------
    constexpr uint32_t IDESC = idesc_i8(SB, NB);
    uint32_t it = 0, tl = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++tl) {
      int j, sb, cb;
      locate(t, j, sb, cb);
      const int nks = jobs[j].nks;
      bar_wait(tempty, (tl & 1) ^ 1);   // the epilogue has drained the accumulators of the previous tile
      tc_fence_after();
      for (int ks = 0; ks < nks; ++ks, ++it) {
        const int st = it % G_::STAGES;
        bar_wait(&full[st], (it / G_::STAGES) & 1);
        tc_fence_after();
        if (lane == 0) {
          const uint32_t abase = s32(smem + (size_t)st * G_::STAGE);
          const uint32_t bbase = abase + G_::A_STEP;
          const uint32_t first = (ks == 0) ? 0u : 1u;
          if (ORDER == 0) {
#pragma unroll
            for (int d = 0; d < NS; ++d) {
#pragma unroll
              for (int i = 0; i <= d; ++i) {
                const uint64_t ad = smem_desc(abase + i * (2 * SB * 16), SB * 16, 128);
                const uint64_t bd = smem_desc(bbase + (d - i) * (2 * NB * 16), NB * 16, 128);
                umma_i8(tmem_base + d * NB, ad, bd, IDESC, (i == 0) ? first : 1u);
              }
            }
          } else {
#pragma unroll
            for (int i = 0; i < NS; ++i) {
              const uint64_t ad = smem_desc(abase + i * (2 * SB * 16), SB * 16, 128);
#pragma unroll
              for (int jj = 0; jj + i < NS; ++jj) {
                const uint64_t bd = smem_desc(bbase + jj * (2 * NB * 16), NB * 16, 128);
                umma_i8(tmem_base + (i + jj) * NB, ad, bd, IDESC, (i == 0) ? first : 1u);
              }
            }
          }
          umma_commit(&empty[st]);             // frees the stage when these MMAs have read it
          if (ks == nks - 1) umma_commit(tfull);  // accumulators complete
        }
        __syncwarp();
      }
    }
  } else {
    // ---------------- epilogue: TMEM -> registers -> fp64 recombination -> G ----------------
    const int quarter = warp & 3;             // TMEM lanes [32 q, 32 q + 32) are this warp's
    const int row = quarter * 32 + lane;      // tile row = TMEM lane
    const uint32_t tlane = tmem_base + ((uint32_t)(quarter * 32) << 16);
    uint32_t tl = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++tl) {
      int j, sb, cb;
      locate(t, j, sb, cb);
      const SliceJob job = jobs[j];
      bar_wait(tfull, tl & 1);
      tc_fence_after();
      double* grow = job.G + ((int64_t)sb * SB + row) * job.ldg + cb * NB;
      const double* cs = job.colscale + cb * NB;
#pragma unroll 1
      for (int c8 = 0; c8 < NB / 8; ++c8) {
        uint32_t r[NS][8];
#pragma unroll
        for (int d = 0; d < NS; ++d) tmem_ld8(tlane + d * NB + c8 * 8, r[d]);
        tmem_ld_wait();
        double out[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          double v = i32_to_f64(r[NS - 1][c]);
#pragma unroll
          for (int d = NS - 2; d >= 0; --d) v = fma(v, 0.00390625, i32_to_f64(r[d][c]));
          out[c] = v * cs[c8 * 8 + c];
        }
#pragma unroll
        for (int c = 0; c < 8; c += 2) *reinterpret_cast<double2*>(grow + c8 * 8 + c) = make_double2(out[c], out[c + 1]);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) bar_arrive(tempty);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(G_::TMEM_COLS) : "memory");
  }
}

// ------------------------------------------------------------------------------------------------------------
// Slicers for operands that exist as fp64 (the table side always; the sample side in the probe and on the paths
// where K1 does not emit slices itself).
// ------------------------------------------------------------------------------------------------------------
// A side: X[r][p] row-major fp64 (ld = ldx), per-pixel inverse scale ic[p] (a power of two; X ic in [-2, 2]).
// One thread per (row, 16-pixel chunk).  rows multiple of 128, npix multiple of 32.
template <int NS>
__global__ void slice_rows_kernel(const double* __restrict__ X, int64_t ldx, const double* __restrict__ ic, int rows, int npix,
                                  int8_t* __restrict__ planes) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int nch = npix / 16;
  if (idx >= (int64_t)rows * nch) return;
  const int r = (int)(idx % rows), ch = (int)(idx / rows);   // consecutive threads = consecutive rows (coalesced pieces)
  const int sb = r / SB, rr = r % SB, ks = ch / 2, c = ch % 2;
  uint64_t dig[16];
#pragma unroll
  for (int p = 0; p < 16; ++p) {
    const int px = ch * 16 + p;
    dig[p] = balanced_digits<NS>(to_fixed<NS>(X[(int64_t)r * ldx + px] * ic[px]));
  }
  int8_t* base = planes + ((int64_t)sb * (npix / KS) + ks) * Geo<NS>::A_STEP + (int64_t)c * (SB * 16) + rr * 16;
#pragma unroll
  for (int i = 0; i < NS; ++i) {
    uint32_t w[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint32_t v = 0;
#pragma unroll
      for (int b = 0; b < 4; ++b) v |= (uint32_t)((dig[4 * q + b] >> (8 * (NS - 1 - i))) & 0xffu) << (8 * b);
      w[q] = v;
    }
    *reinterpret_cast<uint4*>(base + (int64_t)i * (2 * SB * 16)) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

// B side: table T[p][c] row-major fp64 (ld = ldt, ncol columns, zero beyond), per-pixel scale cpix[p] (power of two).
// Pass 1: column scales tau_c = 2^ceil(log2 max_p |T cpix|) (1 for an all-zero column; NaN if any entry is NaN).
__global__ void column_scale_kernel(const double* __restrict__ T, int64_t ldt, const double* __restrict__ cpix, int npix, int ncol,
                                    double* __restrict__ tau, double* __restrict__ colscale /* tau 2^-12 */) {
  const int c = blockIdx.x;
  __shared__ double red[256];
  double m = 0.0;
  bool bad = false;
  for (int p = threadIdx.x; p < npix; p += blockDim.x) {
    const double v = fabs(T[(int64_t)p * ldt + c] * cpix[p]);
    bad = bad || !(v == v);
    m = fmax(m, v);
  }
  red[threadIdx.x] = bad ? __longlong_as_double(0x7ff8000000000000ll) : m;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      const double a = red[threadIdx.x], b = red[threadIdx.x + o];
      red[threadIdx.x] = (a != a || b != b) ? __longlong_as_double(0x7ff8000000000000ll) : fmax(a, b);
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const double mm = red[0];
    double t = 1.0;
    if (mm != mm) t = mm;
    else if (mm > 0.0) {
      int e;
      const double f = frexp(mm, &e);   // mm = f 2^e, f in [0.5, 1)
      t = ldexp(1.0, (f == 0.5) ? e - 1 : e);
      if (!(t < 1.7e308)) t = mm;      // inf
    }
    tau[c] = t;
    colscale[c] = t * 0.000244140625;
  }
}
// Pass 2: planes [cb][K step][slice][2][64][16].  One thread per (column, 16-pixel chunk).
template <int NS>
__global__ void slice_table_kernel(const double* __restrict__ T, int64_t ldt, const double* __restrict__ cpix,
                                   const double* __restrict__ tau, int npix, int ncol_pad, int8_t* __restrict__ planes) {
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int nch = npix / 16;
  if (idx >= (int64_t)ncol_pad * nch) return;
  const int col = (int)(idx % ncol_pad), ch = (int)(idx / ncol_pad);
  const int cb = col / NB, cc = col % NB, ks = ch / 2, c = ch % 2;
  const double it = 1.0 / tau[col];   // power of two (or NaN -> zeros: the NaN travels in colscale)
  uint64_t dig[16];
#pragma unroll
  for (int p = 0; p < 16; ++p) {
    const int px = ch * 16 + p;
    const double v = T[(int64_t)px * ldt + col] * cpix[px] * it;
    dig[p] = balanced_digits<NS>(to_fixed<NS>(v == v ? v : 0.0));
  }
  int8_t* base = planes + ((int64_t)cb * (npix / KS) + ks) * Geo<NS>::B_STEP + (int64_t)c * (NB * 16) + cc * 16;
#pragma unroll
  for (int i = 0; i < NS; ++i) {
    uint32_t w[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint32_t v = 0;
#pragma unroll
      for (int b = 0; b < 4; ++b) v |= (uint32_t)((dig[4 * q + b] >> (8 * (NS - 1 - i))) & 0xffu) << (8 * b);
      w[q] = v;
    }
    *reinterpret_cast<uint4*>(base + (int64_t)i * (2 * NB * 16)) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

}  // namespace sl
}  // namespace dla
#endif
