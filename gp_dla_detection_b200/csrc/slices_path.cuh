// slices_path.cuh -- the product path of the exact-slice int8 contraction (primitives: slices.cuh).
//
// For three lines and the as-written gather (any rank k <= 64), the Woodbury contraction of log_mvnpdf_low_rank.m:22 and
// :26-28 leaves the FP64 pipe: the Voigt kernel emits W and U directly as seven signed 8-bit slices per value
// (six bytes per value instead of the eight of fp64), the per-quasar table [Phi | M] is sliced once, and the products run on the
// integer tcgen05 path with int32 TMEM accumulators, one per slice diagonal.  Pipeline of a chunk of quasars:
//
//   pixel_scales_kernel      per-pixel power-of-two bounds c_w >= max w, c_u >= 2 max |u|          (K0s)
//   table_colscale_kernel    column scales tau of the scaled table [Phi c_w | M c_u]               (K0s)
//   table_slices_kernel      int8 planes of the table, UMMA K-major canonical layout               (K0s)
//   voigt_slices_lane_kernel voigt.c:278-299 + process_qsos.m:189-193 -> int8 planes of W / c_w, U / c_u,
//                            partial sums of r'D^-1 r and log|D| (fp64)                            (K1s)
//   contract_slices_tiles_kernel   G = [W | U] . [Phi | M] on kind::i8 tcgen05, fp64 recombination  (K2s)
//   chol_rows_kernel         per-sample bordered 21 x 21 Cholesky (registers, one quad per sample)  (K2c, k <= 20)
//   chol_generic_kernel      the same for 20 < k <= 64 from the rows of G (kernels.cuh: shared memory, warp per sample)
//
// Fixed point.  Sample side: t = w / c_w in [0, 1) or t = u / c_u in (-1/2, 1/2) enters a double's mantissa by ONE
// fma with a magic constant (16.0 / 24.0): X = round(t 2^48), two's complement, in mantissa bits 0..47.  Its six bytes
// ARE the slices: X = sum_i a_i 256^(5-i), a_1..a_5 plain unsigned bytes, a_0 the top byte -- unsigned for W (t >= 0),
// signed for U; the MMAs of slice 0 of a U unit declare a signed A operand, all others an unsigned one.  So
// t = sum_i a_i 2^(-8 i - 8) with no digit arithmetic at all, rounded at 2^-49 of the pixel's power-of-two scale.
// (-DDLA_NSA=7: seven slices, the bytes of 16 round(t 2^52); 16 x finer, measured 5 % slower end to end.)
// Table side: b = entry / tau in [-1, 1] with 54 fractional bits in seven balanced digits b_j in [-128, 127],
// b = sum_j b_j 2^(-8 j - 6).  Slice products with i + j = d share the weight 2^(-8 d - 14) and accumulate exactly
// in int32 (|sum| <= 7 * 2^15 * pixels < 2^31 for pixels <= 9 000); pairs with i + j >= 7 are dropped: at most
// 6 * 2^-55 of the product of the two scales per pixel -- far below the rounding of the sample value.
#ifndef GPDLA_SLICES_PATH_CUH
#define GPDLA_SLICES_PATH_CUH

#include "slices.cuh"

namespace dla {
namespace sl {

constexpr int NSL = 7;                         // slices per table value = slice diagonals kept
#ifndef DLA_NSA
#define DLA_NSA 6
#endif
// Slices per sample value.  6 (default): X = round(t 2^48), its six bytes are the slices as they lie (sample values
// rounded at 2^-49 of their power-of-two scale; 27 slice pairs).  7: X = round(t 2^52), the bytes of 16 X (28 pairs):
// 16 x finer, +14 % plane bytes, +8 % MMA time (tuning build -DDLA_NSA=7; DESIGN.md 4b has both measured).
constexpr int NSA = DLA_NSA;
static_assert(NSA == 6 || NSA == 7, "sample slices: 6 or 7");
constexpr int A_STEP7 = NSA * 2 * SB * 16;     // bytes of sample planes per K step (32 pixels, 128 samples): 24 576 (28 672)
constexpr int B_STEP64 = NSL * 2 * 64 * 16;    // 14 336 bytes of table planes per K step, 64-column unit
constexpr int B_STEP32 = NSL * 2 * 32 * 16;    //  7 168, 32-column unit
constexpr int GCOLS = 256;                     // row of G: [0, 224) W.Phi (k (k+1) / 2 used), [224, 256) U.M (k used)
constexpr int VOFF_S = 224;
constexpr int UNITS = 5;                       // column units per sample block: 64, 64, 64, 32 (W planes), 32 (U planes)
constexpr int B_COLSTEP = NSL * 2 * 16;        // 224 table bytes per K step and column
// Column layout of a row of G for rank k.  A segment of C columns (C = k (k + 1) / 2 for W.Phi, C = k for U.M) is cut
// into floor(C / 64) units of 64 columns and, for the rest r, one more unit of 64 (r > 32) or 32 (r > 0) columns.
// k <= 20 keeps the fixed layout above (64, 64, 64, 32 | 32: what chol_rows_kernel reads); k = 40: 13 x 64 | 64.
// Unit u starts at column col0(u) of the row AND at byte nks * 224 * col0(u) of the quasar's table planes.
struct SliceLayout {
  int k, T;        // rank, k (k + 1) / 2
  int voff;        // first U.M column = padded width of the W.Phi segment
  int gcols;       // row length of G, tau and colscale
  int nunits, nwu; // units per sample block; the first nwu read the W planes
  int w64, u64;    // 64-column units of the two segments (the 32-column one, if any, is the segment's last)
};
__host__ __device__ inline SliceLayout slice_layout(int k) {
  SliceLayout L;
  L.k = k;
  L.T = k * (k + 1) / 2;
  if (k <= 20) {
    L.voff = VOFF_S; L.gcols = GCOLS; L.nunits = UNITS; L.nwu = 4; L.w64 = 3; L.u64 = 0;
    return L;
  }
  int r = L.T % 64;
  L.w64 = L.T / 64 + (r > 32 ? 1 : 0);
  if (r > 32) r = 0;
  L.nwu = L.w64 + (r > 0 ? 1 : 0);
  L.voff = 64 * L.w64 + (r > 0 ? 32 : 0);
  r = k % 64;
  L.u64 = k / 64 + (r > 32 ? 1 : 0);
  if (r > 32) r = 0;
  L.nunits = L.nwu + L.u64 + (r > 0 ? 1 : 0);
  L.gcols = L.voff + 64 * L.u64 + (r > 0 ? 32 : 0);
  return L;
}
// unit that holds column c: its first column and width
__host__ __device__ inline void unit_of_col(const SliceLayout& L, int c, int& col0, int& N) {
  if (c < L.voff) {
    if (c < 64 * L.w64) { col0 = c & ~63; N = 64; } else { col0 = 64 * L.w64; N = 32; }
  } else {
    const int cu = c - L.voff;
    if (cu < 64 * L.u64) { col0 = L.voff + (cu & ~63); N = 64; } else { col0 = L.voff + 64 * L.u64; N = 32; }
  }
}
constexpr int K2S_STAGE = A_STEP7 + B_STEP64;  // 38 912 (43 008)
// ring depth of the contraction: 5 when it has the SM to itself, 3 (130 KB) when Voigt CTAs of the next chunk share it
constexpr size_t k2s_smem(int stages) { return (size_t)stages * K2S_STAGE + 1024 + 256; }
constexpr int K2S_THREADS = 320;               // warp 0 TMA producer, warp 1 MMA issue, warps 2..9 epilogue (two per TMEM lane quarter)
constexpr int K1Q_THREADS = 128;               // one CTA = the 128 hypotheses of a sample block x one spectrum segment

// CTA residency trace (diagnosis, DLA_B200_TRACE=<file>): {kernel id, SM id, start ns, end ns} per CTA
__device__ unsigned long long* g_trace = nullptr;
__device__ unsigned int g_trace_n = 0;
constexpr unsigned int TRACE_CAP = 1u << 20;
__device__ __forceinline__ unsigned long long trace_now() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void trace_put(unsigned long long kernel_id, unsigned long long t0) {
  unsigned int smid;
  asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
  const unsigned int i = atomicAdd(&g_trace_n, 1u);
  if (i < TRACE_CAP) {
    g_trace[4 * (size_t)i] = kernel_id;
    g_trace[4 * (size_t)i + 1] = smid;
    g_trace[4 * (size_t)i + 2] = t0;
    g_trace[4 * (size_t)i + 3] = trace_now();
  }
}

// per quasar of the staged batch (device array parallel to QuasarDesc)
struct SliceQuasar {
  int64_t a_off;   // byte offset of this quasar's sample planes in the chunk's W (and U) plane buffers
  int64_t b_off;   // byte offset of its table planes
  int32_t nks;     // K steps = ceil(n / 32)
  int32_t pad_;
};

// ------------------------------------------------------------------------------------------------------------
// K0s.  Per-pixel scales.  w = a^2 / (omega2 a^2 + nu) <= 1 / (omega2 + nu) for a in [0, 1];
// |u| = a |y - mu a| / (omega2 a^2 + nu) <= (|y| + |mu|) max_a a / (omega2 a^2 + nu), the maximum at
// a = sqrt(nu / omega2) when that is < 1.  Powers of two, so that scaling is exact.
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ double pow2_above(double x, int extra) {
  if (!(x > 0.0) || !(x < 1.7e308)) return 1.0;   // 0, negative, NaN, inf: the value itself is already meaningless
  int e;
  const double f = frexp(x, &e);                   // x = f 2^e, f in [0.5, 1)
  return ldexp(1.0, e + extra + (f > 0.98 ? 1 : 0));
}
__global__ void pixel_scales_kernel(const QuasarDesc* __restrict__ desc, int q0, const double* __restrict__ flux,
                                    const double* __restrict__ mu, const double* __restrict__ omega2,
                                    const double* __restrict__ noise_variance, double2* __restrict__ isc /* {1/c_w, 1/c_u} */,
                                    double2* __restrict__ csc /* {c_w, c_u} */) {
  const QuasarDesc d = desc[q0 + blockIdx.y];
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= d.n) return;
  const int64_t g = d.pix_off + j;
  const double om = omega2[g], nv = noise_variance[g];
  const double top = 1.0 / (om + nv);
  const double amax = (nv < om) ? 0.5 / sqrt(om * nv) : top;
  const double cw = pow2_above(top, 0);                              // t = w / c_w <= 0.98
  const double cu = pow2_above((fabs(flux[g]) + fabs(mu[g])) * amax, 1);   // |t| = |u| / c_u <= 0.49
  isc[g] = make_double2(1.0 / cw, 1.0 / cu);
  csc[g] = make_double2(cw, cu);
}

// table entry of column c (0 <= c < 256) at pixel p, scaled by the pixel's power of two
__device__ __forceinline__ void col_to_il(int c, const SliceLayout& L, int& i, int& l) {
  const int k = L.k;
  i = -1;
  l = -1;
  if (c < L.T) {
    i = (int)((sqrt(8.0 * c + 1.0) - 1.0) * 0.5);
    while (i * (i + 1) / 2 > c) --i;
    while ((i + 1) * (i + 2) / 2 <= c) ++i;
    l = c - i * (i + 1) / 2;
  } else if (c >= L.voff && c < L.voff + k) {
    i = c - L.voff;
  }
}
__device__ __forceinline__ double table_entry(const double* __restrict__ Mq, int n, int p, int i, int l, double2 cs) {
  const double mi = Mq[(int64_t)i * n + p];
  return (l >= 0) ? (mi * Mq[(int64_t)l * n + p]) * cs.x : mi * cs.y;
}
// column scales: tau_c = 2^ceil(log2 max_p |entry|) (1 for an empty column, NaN if any entry is NaN: a NaN model
// entry must poison every sample of the quasar as it does in fp64); colscale = tau 2^-14
__global__ void table_colscale_kernel(const QuasarDesc* __restrict__ desc, int q0, const double* __restrict__ M, SliceLayout L,
                                      const double2* __restrict__ csc, double* __restrict__ tau,
                                      double* __restrict__ colscale) {
  const QuasarDesc d = desc[q0 + blockIdx.y];
  const int c = blockIdx.x, k = L.k;
  int i, l;
  col_to_il(c, L, i, l);
  __shared__ double red[128];
  double m = 0.0;
  bool bad = false;
  if (i >= 0) {
    const double* Mq = M + (int64_t)k * d.pix_off;
    for (int p = threadIdx.x; p < d.n; p += blockDim.x) {
      const double v = fabs(table_entry(Mq, d.n, p, i, l, csc[d.pix_off + p]));
      bad = bad || !(v == v);
      m = fmax(m, v);
    }
  }
  const double qnan = __longlong_as_double(0x7ff8000000000000ll);
  red[threadIdx.x] = bad ? qnan : m;
  __syncthreads();
  for (int o = blockDim.x / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      const double a = red[threadIdx.x], b = red[threadIdx.x + o];
      red[threadIdx.x] = (a != a || b != b) ? qnan : fmax(a, b);
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const double mm = red[0];
    double t = 1.0;
    if (mm != mm) t = mm;
    else if (mm > 0.0) {
      int e;
      const double f = frexp(mm, &e);
      t = ldexp(1.0, (f == 0.5) ? e - 1 : e);
      if (!(t < 1.7e308)) t = qnan;   // an infinite entry: poison
    }
    tau[(int64_t)blockIdx.y * L.gcols + c] = t;
    colscale[(int64_t)blockIdx.y * L.gcols + c] = t * 6.103515625e-05;
  }
}
// planes of a unit of N columns starting at column col0: [K step][2][slice][N columns][16 B] at b_off + nks 224 col0
// (k <= 20: units 0..2 at u nks B_STEP64, units 3 and 4 of 32 columns behind them).  Within a 16-byte K chunk the slices are stacked like extra columns
// (row j N + c of the operand = slice j, column c), so that ONE MMA multiplies a sample slice with several table
// slices at once.  One thread per (column, 16-pixel chunk); pixels >= n are zero.
__global__ void table_slices_kernel(const QuasarDesc* __restrict__ desc, const SliceQuasar* __restrict__ sq, int q0,
                                    const double* __restrict__ M, SliceLayout L, const double2* __restrict__ csc,
                                    const double* __restrict__ tau, int8_t* __restrict__ planes) {
  const QuasarDesc d = desc[q0 + blockIdx.y];
  const SliceQuasar s = sq[q0 + blockIdx.y];
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const int nch = s.nks * 2, k = L.k;
  if (idx >= L.gcols * nch) return;
  const int col = idx % L.gcols, ch = idx / L.gcols;   // consecutive threads = consecutive columns (contiguous 16-byte pieces)
  int i, l;
  col_to_il(col, L, i, l);
  const double tc = tau[(int64_t)blockIdx.y * L.gcols + col];
  const double it = 1.0 / tc;   // a power of two; NaN -> zeros here, the NaN travels in colscale
  const double* Mq = M + (int64_t)k * d.pix_off;
  uint64_t dig[16];
#pragma unroll
  for (int p = 0; p < 16; ++p) {
    const int px = ch * 16 + p;
    double v = 0.0;
    if (i >= 0 && px < d.n) v = table_entry(Mq, d.n, px, i, l, csc[d.pix_off + px]) * it;
    dig[p] = balanced_digits<NSL>(to_fixed<NSL>(v == v ? v : 0.0));
  }
  int col0, N;
  unit_of_col(L, col, col0, N);
  const int cc = col - col0;
  const int64_t ubase = (int64_t)s.nks * B_COLSTEP * col0;
  const int ks = ch >> 1, c2 = ch & 1;
  int8_t* base = planes + s.b_off + ubase + (int64_t)ks * (NSL * 2 * N * 16) + (int64_t)c2 * (NSL * N * 16) + cc * 16;
#pragma unroll
  for (int sl_ = 0; sl_ < NSL; ++sl_) {
    uint32_t w[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      uint32_t v = 0;
#pragma unroll
      for (int b = 0; b < 4; ++b) v |= (uint32_t)((dig[4 * q + b] >> (8 * (NSL - 1 - sl_))) & 0xffu) << (8 * b);
      w[q] = v;
    }
    *reinterpret_cast<uint4*>(base + (int64_t)sl_ * (N * 16)) = make_uint4(w[0], w[1], w[2], w[3]);
  }
}

// ------------------------------------------------------------------------------------------------------------
// K1s.  voigt_weights_lane_kernel (kernels.cuh) with the slice emission: a lane owns one hypothesis and walks one
// segment of the spectrum; segments are whole 16-pixel chunks.  Every step yields four pixels, i.e. one 32-bit
// word of each of the 14 sixteen-byte pieces (7 slices of W, 7 of U) the lane owns; the pieces are complete after
// four steps and leave as 16-byte stores, 512 contiguous bytes per warp and piece.  The words of the first three
// steps wait in a per-lane shared-memory slot ([slice quad][step][lane] x 16 B), not in registers.
//   planes: a_off + ((sb * nks + ks) * 7 + slice) * 4096 + (chunk & 1) * 2048 + row * 16
// ------------------------------------------------------------------------------------------------------------
// m[h] = fma(value, 1 / scale, MAGIC): bits(m) - bits(MAGIC) = X = round(t 2^52) (two's complement, |X| < 2^52);
// slice i = byte 6 - i of 16 X; out[i] = the four pixels' slice-i bytes.
// The exponent field of m never reaches a slice (the top slice is bits 16..23 of the shifted high word = mantissa bits
// 44..51), so bits(MAGIC) need not be subtracted: for W (MAGIC = 1.0) the mantissa is X itself; for U (MAGIC = 1.5)
// it is X + 2^51, i.e. the two's-complement X with its sign bit flipped -- one xor on the top-slice word (TOPFLIP).
template <bool TOPFLIP>
__device__ __forceinline__ void digits4(const double (&m)[4], uint32_t* out /* [NSA] */) {
  uint32_t lo[4], hi[4];
#pragma unroll
  for (int h = 0; h < 4; ++h) {
    const uint32_t l = (uint32_t)__double2loint(m[h]), hh = (uint32_t)__double2hiint(m[h]);
    if (NSA == 7) {
      lo[h] = l << 4;                         // 16 X: the top slice becomes bits 44..51 of X
      hi[h] = __funnelshift_l(l, hh, 4);
    } else {
      lo[h] = l;                              // X = round(t 2^48) sits in mantissa bits 0..47 as it is
      hi[h] = hh;
    }
  }
  const uint32_t t0 = __byte_perm(lo[0], lo[1], 0x5140), t1 = __byte_perm(lo[2], lo[3], 0x5140);
  const uint32_t t2 = __byte_perm(lo[0], lo[1], 0x7362), t3 = __byte_perm(lo[2], lo[3], 0x7362);
  out[NSA - 1] = __byte_perm(t0, t1, 0x5410);
  out[NSA - 2] = __byte_perm(t0, t1, 0x7632);
  out[NSA - 3] = __byte_perm(t2, t3, 0x5410);
  out[NSA - 4] = __byte_perm(t2, t3, 0x7632);
  const uint32_t t4 = __byte_perm(hi[0], hi[1], 0x5140), t5 = __byte_perm(hi[2], hi[3], 0x5140);
  out[NSA - 5] = __byte_perm(t4, t5, 0x5410);
  out[NSA - 6] = __byte_perm(t4, t5, 0x7632);
  if (NSA == 7) {
    const uint32_t t6 = __byte_perm(hi[0], hi[1], 0x0062), t7 = __byte_perm(hi[2], hi[3], 0x0062);
    // the mantissa of 1.5 + t is X + 2^51: two's-complement X with its sign bit flipped -- one xor on the top-slice word
    out[0] = __byte_perm(t6, t7, 0x5410) ^ (TOPFLIP ? 0x80808080u : 0u);
  }
}
constexpr double MAGIC_W = (NSA == 7) ? 1.0 : 16.0;   // ulp(MAGIC) = 2^-52 or 2^-48: the fma rounds t to that grid
constexpr double MAGIC_U = (NSA == 7) ? 1.5 : 24.0;   // + 2^51 ulps: negative t stays in the same binade

// MINB: resident CTAs per SM the kernel is compiled for (3 -> 162 registers; 4 -> 126, the default, no spills).  PXS: the per-pixel records
// {omega2, noise, mu, flux, 1/c_w, 1/c_u} of the segment are staged in shared memory (px_cap pixels, 48 B each);
// otherwise they are warp-uniform global loads (L1 hits) and a CTA needs 26 KB only.
template <int MINB, bool PXS>
__global__ void __launch_bounds__(K1Q_THREADS, MINB)
voigt_slices_lane_kernel(const QuasarDesc* __restrict__ desc, const SliceQuasar* __restrict__ sq, int q0, int nsb /* 128-sample blocks */,
                         int64_t S, const double* __restrict__ offset_sorted, const double* __restrict__ nhi_sorted,
                         const double* __restrict__ padded_wavelengths, const double2* __restrict__ pix4,
                         const double2* __restrict__ isc, const double* __restrict__ etab256, int8_t* __restrict__ Wplanes,
                         int8_t* __restrict__ Uplanes, double* __restrict__ aux, int lam_cap, int px_cap, int nseg) {
  extern __shared__ __align__(16) double k1q_smem[];
  __shared__ double etab_s[256];
  const bool tracing = g_trace != nullptr;
  const unsigned long long trace_t0 = (tracing && threadIdx.x == 0) ? trace_now() : 0ull;
  const int seg = blockIdx.x % nseg;
  const int tile = blockIdx.x / nseg;
  const int ql = tile / nsb, sb = tile % nsb;
  const QuasarDesc d = desc[q0 + ql];
  const SliceQuasar sqq = sq[q0 + ql];
  const int tid = threadIdx.x;
  const int n = d.n, Pm1 = d.P - 1;
  const int nch = 2 * sqq.nks;                                            // 16-pixel chunks
  const int c0 = (int)(((int64_t)seg * nch) / nseg), c1 = (int)(((int64_t)(seg + 1) * nch) / nseg);
  const int g0 = 4 * c0, g1 = 4 * c1;                                     // groups of four pixels
  const int lbase = max(4 * g0 - 2, 0), llast = min(4 * g1 + 5, Pm1);
  const int j0 = min(4 * g0, n), j1 = min(4 * g1, n);                     // kept pixels this segment emits
  const int nsteps = (g1 > g0) ? g1 - g0 + 2 : 0;
  double* lam_s = k1q_smem;
  double2* px_s = reinterpret_cast<double2*>(k1q_smem + lam_cap);
  // PXS: 48-byte records {omega2, noise}, {mu, flux}, {1/c_w, 1/c_u} per pixel
  uint4* stg_s = reinterpret_cast<uint4*>(px_s + 3 * (PXS ? px_cap : 0)) + tid;   // [4 slice quads][3 steps][128 lanes] x 16 B
  for (int i = tid; i < 256; i += K1Q_THREADS) etab_s[i] = etab256[i];
  {
    const double* lam_g = padded_wavelengths + d.pad_off;
    for (int i = tid; i < 4 * nsteps; i += K1Q_THREADS) lam_s[i] = __ldg(lam_g + min(max(4 * g0 - 2 + i, lbase), llast));
    if (PXS) {
      const double2* px_g = pix4 + 2 * (d.pix_off + j0);
      const double2* sc_g = isc + d.pix_off + j0;
      for (int i = tid; i < 3 * (j1 - j0); i += K1Q_THREADS) {
        const int pxl = i / 3, part = i - 3 * pxl;
        px_s[i] = (part < 2) ? __ldg(px_g + 2 * pxl + part) : __ldg(sc_g + pxl);
      }
    }
  }
  __syncthreads();
  const int64_t s = (int64_t)sb * K1Q_THREADS + tid;
  const bool real = s < S;
  const double off = real ? offset_sorted[s] : 0.0;       // process_qsos.m:163-165, separately rounded
  const double N = real ? nhi_sorted[s] : 0.0;
  const double z = __dadd_rn(d.min_z, __dmul_rn(__dsub_rn(d.max_z, d.min_z), off));
  const double opz = __dadd_rn(1.0, z);
  double mult[3], ncw[3], nc1[3], nc2[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    mult[j] = DLA_C_CGS / __dmul_rn(dla_transition_wavelengths[j], opz) / 1e8;  // voigt.c:278-279
    ncw[j] = -N * g_lines.cwT[j];
    nc1[j] = ncw[j] * g_lines.g1p[j];
    nc2[j] = ncw[j] * g_lines.gq[2];
  }
  constexpr int PSTR = PXS ? 3 : 2;                                     // double2's per pixel record
  const double2* px = PXS ? px_s - 3 * j0 : pix4 + 2 * d.pix_off;      // indexed by PSTR * kept pixel
  const double2* sc = isc + d.pix_off;                                  // !PXS: the reciprocal scales, by kept pixel
  const uint32_t lam_a = smem_u32(lam_s);                               // 32-bit shared-window addresses: no generic
  const SmemTab etab{smem_u32(etab_s)};                                 // pointer arithmetic inside the step loop
  const int jdead = (j1 > j0) ? j0 : 0;  // a live pixel of this segment stands in for dead ones (values discarded)
  const int gfull = n >> 2;
  const int64_t S_pad = (int64_t)nsb * SB;
  const int row = tid;
  const int64_t sb_base = sqq.a_off + (int64_t)sb * sqq.nks * A_STEP7 + row * 16;
  int8_t* wq = Wplanes + sb_base;
  int8_t* uq = Uplanes + sb_base;

  double q0acc = 0.0, ldacc = 0.0, dprod = 1.0;
  int esum = 0;
  double win[10];
#pragma unroll
  for (int t = 0; t < 10; ++t) win[t] = 0.0;
  const bool have_px = j1 > j0;   // a segment wholly beyond n (padding chunks) emits zeros
  for (int st = 0; st < nsteps; ++st) {
    double l4[4], r4[4];
    {
      const double2 la = lds_f64x2_ro(lam_a + 32u * (uint32_t)st), lb = lds_f64x2_ro(lam_a + 32u * (uint32_t)st + 16u);
      l4[0] = la.x, l4[1] = la.y, l4[2] = lb.x, l4[3] = lb.y;
    }
    raw_lane<4, SmemTab>(l4, mult, ncw, nc1, nc2, N, etab, r4);
#pragma unroll
    for (int t = 0; t < 6; ++t) win[t] = win[t + 4];
#pragma unroll
    for (int h = 0; h < 4; ++h) win[6 + h] = r4[h];
    if (st >= 2) {
      const int g = g0 + st - 2;
      double w[4], u[4];
      double2 is[4];
      if (g < gfull) {
        weights_lane_core<false, PSTR>(win, g, n, jdead, px, w, u, q0acc, ldacc, dprod, esum, is);
        if (!PXS) {
#pragma unroll
          for (int h = 0; h < 4; ++h) is[h] = sc[4 * g + h];
        }
      } else if (have_px) {
        weights_lane_core<true, PSTR>(win, g, n, jdead, px, w, u, q0acc, ldacc, dprod, esum, is);   // w = u = 0 beyond n
        if (!PXS) {
#pragma unroll
          for (int h = 0; h < 4; ++h) is[h] = sc[(4 * g + h < n) ? 4 * g + h : jdead];
        }
      } else {
#pragma unroll
        for (int h = 0; h < 4; ++h) w[h] = u[h] = 0.0, is[h] = make_double2(0.0, 0.0);
      }
      double mw[4], mu_[4];
#pragma unroll
      for (int h = 0; h < 4; ++h) {
        mw[h] = fma(w[h], is[h].x, MAGIC_W);
        mu_[h] = fma(u[h], is[h].y, MAGIC_U);
      }
      uint32_t nw[NSA + 2], nu[NSA + 2];   // padded to two quads
      nw[NSA] = nw[NSA + 1] = nu[NSA] = nu[NSA + 1] = 0u;
      digits4<false>(mw, nw);
      digits4<true>(mu_, nu);
      const int qs = g & 3;
      if (qs < 3) {
        // slice quads {W0..W3}, {W4.., 0}, {U0..U3}, {U4.., 0}: four 16-byte stores per step instead of 2 NSA words
        uint4* slot = stg_s + qs * K1Q_THREADS;
        slot[0] = make_uint4(nw[0], nw[1], nw[2], nw[3]);
        slot[3 * K1Q_THREADS] = make_uint4(nw[4], nw[5], nw[6], nw[7]);
        slot[6 * K1Q_THREADS] = make_uint4(nu[0], nu[1], nu[2], nu[3]);
        slot[9 * K1Q_THREADS] = make_uint4(nu[4], nu[5], nu[6], nu[7]);
      } else {
        const int ch = g >> 2;
        const int64_t o = (int64_t)(ch >> 1) * A_STEP7 + (ch & 1) * (SB * 16);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint4 s0 = stg_s[(3 * q) * K1Q_THREADS], s1 = stg_s[(3 * q + 1) * K1Q_THREADS], s2 = stg_s[(3 * q + 2) * K1Q_THREADS];
          constexpr int PL = 2 * SB * 16;   // bytes of one slice plane of a K step
          int8_t* dst = ((q < 2) ? wq : uq) + o + (q & 1) * 4 * PL;
          const uint32_t* last = ((q < 2) ? nw : nu) + (q & 1) * 4;
          const int sl0 = (q & 1) * 4;
          *reinterpret_cast<uint4*>(dst) = make_uint4(s0.x, s1.x, s2.x, last[0]);
          *reinterpret_cast<uint4*>(dst + PL) = make_uint4(s0.y, s1.y, s2.y, last[1]);
          if (sl0 + 2 < NSA) *reinterpret_cast<uint4*>(dst + 2 * PL) = make_uint4(s0.z, s1.z, s2.z, last[2]);
          if (sl0 + 3 < NSA) *reinterpret_cast<uint4*>(dst + 3 * PL) = make_uint4(s0.w, s1.w, s2.w, last[3]);
        }
      }
    }
  }
  ldacc += fma((double)esum, 0.693147180559945309417232, log(dprod));
  double* auxq = aux + (int64_t)ql * 2 * nseg * S_pad;
  auxq[(int64_t)seg * S_pad + s] = q0acc;
  auxq[(int64_t)(nseg + seg) * S_pad + s] = ldacc;
  if (tracing && threadIdx.x == 0) trace_put(1, trace_t0);
}

// ------------------------------------------------------------------------------------------------------------
// K2s.  Persistent: CTA b walks the work items t = b, b + gridDim.x, ...; item t = ((ql * nsb) + sb) * 5 + unit, so the
// units of a sample block (five at k <= 20) run at the same time on neighbouring SMs and share the block's planes
// through L2.  Warp 0: TMA producer (two 1-D bulk copies per K step), warp 1: MMA issue (10 or 7 wide kind::i8 MMAs per
// K step), warps 2..9: epilogue (TMEM -> two register buffers -> fp64 -> G).  MODE 1 / 2 / 3 (diagnosis): no MMA / no loads after the first ring fill / no epilogue work.
// ------------------------------------------------------------------------------------------------------------
struct K2sTile {
  const int8_t* a;
  const int8_t* b;
  int nks, N, ql, sb, col0;
  uint32_t isu;   // the unit reads the U planes (top slice signed)
};
__device__ __forceinline__ K2sTile k2s_decode(int t, const SliceQuasar* __restrict__ sq, int q0, int nsb, const SliceLayout& L,
                                              const int8_t* Aw, const int8_t* Au, const int8_t* Bt) {
  K2sTile r;
  const int per = nsb * L.nunits;
  r.ql = t / per;
  const int rem = t - r.ql * per;
  r.sb = rem / L.nunits;
  const int u = rem - r.sb * L.nunits;
  const SliceQuasar s = sq[q0 + r.ql];
  r.nks = s.nks;
  const bool isu = u >= L.nwu;
  const int uu = isu ? u - L.nwu : u, n64 = isu ? L.u64 : L.w64;
  r.N = (uu < n64) ? 64 : 32;
  r.isu = isu ? 1u : 0u;
  r.col0 = (isu ? L.voff : 0) + 64 * ((uu < n64) ? uu : n64);
  r.a = (isu ? Au : Aw) + s.a_off + (int64_t)r.sb * s.nks * A_STEP7;
  r.b = Bt + s.b_off + (int64_t)s.nks * B_COLSTEP * r.col0;
  return r;
}

// One K step of a tile: sample slice i meets the table slices 0 .. 6 - i in ONE wide MMA (two when that is more than
// 256 columns): operand rows [0, (7 - i) N) of the stacked table planes, accumulators of the diagonals i .. 6.
// 10 MMAs (N = 64) or 7 (N = 32) instead of 28: the sample slice is read from shared memory once per MMA, and the
// single issuing thread is no longer the limit (a 128 x 64 x 32 MMA lasts 32 clocks, issuing one costs about 70).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
// a 2^16 + b 2^8 + c for three int32 accumulators: exact, below 2^48 in magnitude
__device__ __forceinline__ long long combine3(int a, int b, int c) {
  return (((long long)a << 8) + (long long)b << 8) + (long long)c;
}
// int64 (|x| < 2^51) -> double without the conversion unit: x + 2^51 placed in the mantissa of 2^52, minus the bias
__device__ __forceinline__ double i64_to_f64(long long x) {
  return __longlong_as_double((x + 0x0008000000000000ll) + 0x4330000000000000ll) - 6755399441055744.0;   // 2^52 + 2^51
}
template <int N>
__device__ __forceinline__ void k2s_issue_step(uint32_t abase, uint32_t bbase, uint32_t tmem_base, uint32_t first, uint32_t top_signed) {
  constexpr uint64_t HI = (uint64_t)(((128u >> 4) & 0x3FFFu) | (1u << 14)) << 32;   // stride between 8-row groups, version 1
  const uint32_t a_lo = ((abase >> 4) & 0x3FFFu) | ((uint32_t)((SB * 16) >> 4) << 16);
  const uint32_t b_lo = ((bbase >> 4) & 0x3FFFu) | ((uint32_t)((NSL * 16 * N) >> 4) << 16);
#pragma unroll
  for (int i = 0; i < NSA; ++i) {
    const uint64_t ad = HI | (uint64_t)(a_lo + (uint32_t)(i * ((2 * SB * 16) >> 4)));
    const int ntot = (NSL - i) * N;
#pragma unroll
    for (int r0 = 0; r0 < ntot; r0 += 256) {
      const int np = (ntot - r0 < 256) ? ntot - r0 : 256;
      const uint64_t bd = HI | (uint64_t)(b_lo + (uint32_t)r0);   // r0 rows of 16 bytes
      // A format (bits [7, 10)): 1 = signed 8 bit for the top slice of U, 0 = unsigned for W's top slice and the plain low bytes
      umma_i8(tmem_base + (uint32_t)(i * N + r0), ad, bd, (idesc_i8(SB, np) & ~(1u << 7)) | ((i == 0) ? (top_signed << 7) : 0u), (i == 0) ? first : 1u);
    }
  }
}

// one CTA per SM (shared memory): up to 204 registers per thread, the epilogue warps use them for two buffers
template <int MODE, int K2S_STAGES>
__global__ void __launch_bounds__(K2S_THREADS, 1)
contract_slices_tiles_kernel(const SliceQuasar* __restrict__ sq, int q0, int nq, int nsb, const SliceLayout L, const int8_t* __restrict__ Aw,
                             const int8_t* __restrict__ Au, const int8_t* __restrict__ Bt, const double* __restrict__ colscale,
                             double* __restrict__ G) {
  extern __shared__ unsigned char sl_smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(sl_smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)K2S_STAGES * K2S_STAGE);
  uint64_t* empty = full + K2S_STAGES;
  uint64_t* tfull = empty + K2S_STAGES;
  uint64_t* tempty = tfull + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 1);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ntiles = nq * nsb * L.nunits;
  const bool tracing = g_trace != nullptr;
  const unsigned long long trace_t0 = (tracing && tid == 0) ? trace_now() : 0ull;

  if (tid == 0) {
    for (int s = 0; s < K2S_STAGES; ++s) {
      bar_init(&full[s], 1);
      bar_init(&empty[s], 1);
    }
    bar_init(tfull, 1);
    bar_init(tempty, 8);   // one arrive per epilogue warp
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(tmem_slot)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const K2sTile tl = k2s_decode(t, sq, q0, nsb, L, Aw, Au, Bt);
        const uint32_t bstep = (uint32_t)(NSL * 2 * 16) * tl.N;
        for (int ks = 0; ks < tl.nks; ++ks, ++it) {
          const int st = it % K2S_STAGES;
          bar_wait(&empty[st], ((it / K2S_STAGES) & 1) ^ 1);
          unsigned char* dst = smem + (size_t)st * K2S_STAGE;
          if (MODE == 2 && it >= (uint32_t)K2S_STAGES) {
            bar_arrive(&full[st]);
          } else {
            bar_expect_tx(&full[st], A_STEP7 + bstep);
            bulk_load(dst, tl.a + (int64_t)ks * A_STEP7, A_STEP7, &full[st]);
            bulk_load(dst + A_STEP7, tl.b + (int64_t)ks * bstep, bstep, &full[st]);
          }
        }
      }
    }
  } else if (warp == 1) {
    uint32_t it = 0, tcount = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++tcount) {
      const K2sTile tl = k2s_decode(t, sq, q0, nsb, L, Aw, Au, Bt);
      bar_wait(tempty, (tcount & 1) ^ 1);   // the epilogue has drained the previous tile's accumulators
      tc_fence_after();
      for (int ks = 0; ks < tl.nks; ++ks, ++it) {
        const int st = it % K2S_STAGES;
        bar_wait(&full[st], (it / K2S_STAGES) & 1);
        tc_fence_after();
        // every lane forms the (warp-uniform) operands; one elected lane issues
        const uint32_t abase = s32(smem + (size_t)st * K2S_STAGE);
        const uint32_t first = (ks == 0) ? 0u : 1u;
        const bool last = ks == tl.nks - 1;
        if (elect_one()) {
          if (MODE != 1) {
            if (tl.N == 64) k2s_issue_step<64>(abase, abase + A_STEP7, tmem_base, first, tl.isu);
            else k2s_issue_step<32>(abase, abase + A_STEP7, tmem_base, first, tl.isu);
          }
          umma_commit(&empty[st]);
          if (last) umma_commit(tfull);
        }
        __syncwarp();
      }
    }
  } else {
    const int quarter = warp & 3;             // TMEM lanes [32 q, 32 q + 32) belong to the warps with warp % 4 == q
    const int half = (warp - 2) >> 2;         // the two warps of a quarter split the columns
    const int row = quarter * 32 + lane;
    const uint32_t tlane = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const int64_t S_pad = (int64_t)nsb * SB;
    uint32_t tcount = 0;
    for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++tcount) {
      const K2sTile tl = k2s_decode(t, sq, q0, nsb, L, Aw, Au, Bt);
      bar_wait(tfull, tcount & 1);
      tc_fence_after();
      double* grow = G + ((int64_t)tl.ql * S_pad + (int64_t)tl.sb * SB + row) * L.gcols + tl.col0;
      const double* cs = colscale + (int64_t)tl.ql * L.gcols + tl.col0;
      const int nc8h = (MODE == 3) ? 0 : tl.N / 16;   // 4 or 2 groups of eight columns for this warp
      const int cb = half * nc8h;
      // two register buffers: the TMEM loads of group i + 1 are in flight while group i is recombined and stored, and
      // the accumulators are handed back to the MMA warp as soon as the last group sits in registers
      uint32_t r[2][NSL][8];
      auto load = [&](int i, uint32_t (&dst)[NSL][8]) {
#pragma unroll
        for (int dgn = 0; dgn < NSL; ++dgn) tmem_ld8(tlane + (uint32_t)(dgn * tl.N + (cb + i) * 8), dst[dgn]);
      };
      auto emit = [&](int i, const uint32_t (&src)[NSL][8]) {
        const int c8 = cb + i;
        double out[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
#if defined(DLA_K2S_EPI_HORNER)   // tuning build: seven conversions and six fma per column
          double v = i32_to_f64(src[NSL - 1][c]);
#pragma unroll
          for (int dgn = NSL - 2; dgn >= 0; --dgn) v = fma(v, 0.00390625, i32_to_f64(src[dgn][c]));
          out[c] = v * __ldg(cs + c8 * 8 + c);
#else
          // three diagonals at a time, exactly, in 64-bit integers (|.| < 2^48), then two fp64 steps instead of six
          const long long g0 = combine3((int)src[0][c], (int)src[1][c], (int)src[2][c]);
          const long long g1 = combine3((int)src[3][c], (int)src[4][c], (int)src[5][c]);
          double v = fma(i32_to_f64(src[6][c]), 0.00390625, i64_to_f64(g1));             // g1 + acc_6 2^-8
          v = fma(v, 5.9604644775390625e-08, i64_to_f64(g0));                            // g0 + (...) 2^-24
          out[c] = v * (__ldg(cs + c8 * 8 + c) * 1.52587890625e-05);                     // sum_d acc_d 2^(-8 d) = v 2^-16
#endif
        }
#pragma unroll
        for (int c = 0; c < 8; c += 2) *reinterpret_cast<double2*>(grow + c8 * 8 + c) = make_double2(out[c], out[c + 1]);
      };
      if (nc8h > 0) load(0, r[0]);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (i < nc8h) {
          tmem_ld_wait();
          if (i + 1 < nc8h) load(i + 1, r[(i + 1) & 1]);
          else {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) bar_arrive(tempty);
          }
          emit(i, r[i & 1]);
        }
      }
      if (nc8h == 0) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) bar_arrive(tempty);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
  }
  if (tracing && tid == 0) trace_put(2, trace_t0);
}

// ------------------------------------------------------------------------------------------------------------
// K2c.  log_mvnpdf_low_rank.m:23-32 from the rows of G: one quad of lanes per sample (chol_logp_quad, kernels.cuh),
// K2C_ROWS samples per CTA, their rows (2 KB each) brought in by as many bulk copies into padded shared-memory rows.
// ------------------------------------------------------------------------------------------------------------
#ifndef DLA_K2C_ROWS
#define DLA_K2C_ROWS 16
#endif
#ifndef DLA_K2C_MINB
#define DLA_K2C_MINB 5
#endif
constexpr int K2C_ROWS = DLA_K2C_ROWS;   // samples per CTA: 64 threads, five CTAs = 10 warps per SM at 192 registers (32 rows, 2 CTAs: 8 warps, +18 % time)
constexpr int K2C_LD = GCOLS + 2;   // padded row (doubles): the eight samples of a warp start in different banks
constexpr size_t K2C_SMEM = (size_t)K2C_ROWS * K2C_LD * 8 + 16;
__global__ void __launch_bounds__(K2C_ROWS * 4, DLA_K2C_MINB)
chol_rows_kernel(const QuasarDesc* __restrict__ desc, int q0, int nsb /* 128-sample blocks */, int64_t S,
                 const double* __restrict__ G, const double* __restrict__ aux, const int32_t* __restrict__ perm,
                 double* __restrict__ sample_ll /* [nq][S], caller order */, double* __restrict__ ll_no, int k, int nseg) {
  extern __shared__ __align__(16) unsigned char k2c_smem[];
  double* rows = reinterpret_cast<double*>(k2c_smem);
  uint64_t* bar = reinterpret_cast<uint64_t*>(k2c_smem + (size_t)K2C_ROWS * K2C_LD * 8);
  const int ql = blockIdx.y;
  const QuasarDesc d = desc[q0 + ql];
  const int64_t S_pad = (int64_t)nsb * SB;
  const int64_t s0 = (int64_t)blockIdx.x * K2C_ROWS;
  if (s0 > S) return;   // rows beyond the no-DLA pseudo-sample are padding
  const int tid = threadIdx.x;
  if (tid == 0) {
    bar_init(bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (tid == 0) {
    bar_expect_tx(bar, K2C_ROWS * GCOLS * 8);
    const double* src = G + ((int64_t)ql * S_pad + s0) * GCOLS;
    for (int r = 0; r < K2C_ROWS; ++r) bulk_load(rows + (size_t)r * K2C_LD, src + (size_t)r * GCOLS, GCOLS * 8, bar);
  }
  const int srow = tid >> 2, t = tid & 3;
  const int64_t s = s0 + srow;
  double q0v, ldv;
  aux_sums(aux, ql, nseg, S_pad, s, q0v, ldv);
  bar_wait(bar, 0);
  const double lp = chol_logp_quad<0, VOFF_S>(rows + (size_t)srow * K2C_LD, k, t, q0v, ldv, d.n);
  if (t == 0) {
    if (s < S) sample_ll[(int64_t)ql * S + perm[s]] = lp;
    else if (s == S) ll_no[q0 + ql] = lp;
  }
}

}  // namespace sl
}  // namespace dla
#endif
