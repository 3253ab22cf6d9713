// dla_b200.cu -- C ABI (include/dla_b200.h) and host orchestration of the sm_100a kernels.
// No torch types, no CPU fallback: every compute entry point launches CUDA kernels.
#include "../../include/dla_b200.h"

#include <cuda_runtime.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <exception>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include <nvtx3/nvToolsExt.h>   // header-only: ranges are no-ops unless a profiler injects its library

#include "kernels.cuh"
#include "slices_path.cuh"

using dla::QuasarDesc;
using dla::sl::SliceQuasar;

namespace {

thread_local std::string g_create_error;

// NVTX range over a scope: the host side of one stage (copies, K0..K3 launches) of one chunk, as a timeline tool sees it
struct NvtxRange {
  explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
  ~NvtxRange() { nvtxRangePop(); }
};

// DLA_B200_VERBOSE=1: one line per chunk on stderr, like fprintf_debug in process_qsos.m:93-94, :219
bool g_verbose = [] {
  const char* v = getenv("DLA_B200_VERBOSE");
  return v && *v && strcmp(v, "0") != 0;
}();

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e == cudaSuccess) cap = bytes;
    return e;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
  template <typename T>
  T* as() const { return reinterpret_cast<T*>(p); }
};

// grow-only page-locked host buffer (staging for arrays the library builds itself)
struct PinBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t ensure(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
    const size_t want = bytes + bytes / 4;
    cudaError_t e = cudaHostAlloc(&p, want, cudaHostAllocDefault);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() {
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
  }
  template <typename T>
  T* as() const { return reinterpret_cast<T*>(p); }
};

struct Chunk {
  int q0, nq;
  int64_t wu_doubles, bmat_doubles;
  int max_npad, max_P;
  bool all_ident;  // every quasar of the chunk uses the as-written gather (profile_index[j] == j)
  // exact-slice int8 contraction (slices_path.cuh): bytes of one operand's sample planes / of the table planes
  int64_t a_bytes, b_bytes;
  int max_nks;
  bool slices;     // this chunk runs K0s/K1s/K2s/K2c instead of the DMMA kernels
};

// Device-side views of one staged batch.
struct Inputs {
  const QuasarDesc* desc;
  const double *padded, *flux, *mu, *om, *nv, *M;
  const int32_t* pidx;
  double2* pix4;  // packed {omega2, noise, mu, flux} per kept pixel (built per chunk on the device)
  const SliceQuasar* sq = nullptr;  // slice geometry per quasar (parallel to desc)
};

}  // namespace

struct dla_ctx {
  int device = 0;
  dla_params p{};
  cudaStream_t stream = nullptr;
  std::string err;
  cudaStream_t s_in = nullptr, s_out = nullptr;  // copy streams: H2D ahead of, D2H behind, the compute stream
  cudaEvent_t ev[8]{};
  std::vector<cudaEvent_t> ev_time;   // NTEV timing events per chunk
  std::vector<cudaEvent_t> ev_sync;   // 2 ordering events per chunk: inputs resident, kernels done
  std::vector<Chunk> chunk_copy;      // host-side copy of the plan the events belong to
  // host pointers of the batch being staged chunk by chunk (valid during dla_eval_batch only)
  struct HostIn {
    const int64_t *pix_ptr, *pad_ptr;
    const double *padded, *flux, *mu, *om, *nv, *M;
    const int32_t* pidx;
    const double *lamk, *zq;  // device-prep mode: kept wavelengths [sum n_q], z_qso [Q]
  } hin{};
  // learned model for the device-side preparation (dla_set_model)
  bool have_model = false, dev_prep = false;
  int64_t n_rest = 0;
  dla_prep_params prep{};
  DevBuf d_mrest, d_mmu, d_mM, d_mlo, d_lamk, d_zq;
  // dla_eval_spectra's compacted arrays of one sub-batch of quasars (page-locked: async H2D); two sets, so that a
  // host thread can select the next sub-batch while the device works on the current one
  struct SpectraStage {
    PinBuf lamk, fk, nk, padded, pidx;
    std::vector<int64_t> pix_ptr, pad_ptr;
    std::vector<double> minz, maxz;
    int64_t nkept = 0;
    int rc = 0;
    const char* err = nullptr;
  } sstage[2];
  // training objective (dla_train_set_data / dla_train_objective)
  int64_t tQ = 0, tN = 0;
  DevBuf t_y, t_l, t_nv, t_x, t_om2, t_aux, t_E, t_fq, t_beta, t_P, t_A1, t_S, t_gp, t_g, t_gsc, t_desc;
  // hypothesis set
  // (sorted by offset once: the hypotheses of a warp are then neighbours in z; d_perm maps a
  //  sorted position back to the caller's sample index)
  DevBuf d_offset, d_nhi, d_perm, d_etab256;
  bool have_samples = false;
  DevBuf seam[11];  // dla_voigt / dla_log_mvnpdf_low_rank: grow-only device arrays (no cudaMalloc per call)
  double* noalloc_scratch = nullptr;  // device address of g_noalloc_scratch on this context's GPU (dla_voigt_noalloc)
  std::vector<double> h_set_offset, h_set_nhi, h_set_model;  // what the device holds now, as the caller passed it
  bool force_pixel_lane_k1 = false;  // DLA_B200_K1=pixel: always use voigt_weights_kernel
  bool force_tile_k2 = false;        // DLA_B200_K2=tile: one CTA per tile (contract_chol_kernel)
  bool k2_rowsplit = false;          // DLA_B200_K2=rowsplit: contract_chol_rowsplit_kernel
  bool use_fused = false;            // DLA_B200_FUSED=1: fused_voigt_contract_kernel (W/U never leave the SM)
  bool use_slices = true;            // DLA_B200_CONTRACT=dmma turns the int8 tcgen05 contraction off (FP64 DMMA kernels)
  int k2s_mode = 0;                  // DLA_B200_K2S_MODE=1/2: diagnosis builds of the contraction (no MMA / no loads)
  int k1q_variant = 1;               // DLA_B200_K1Q=0/1/2: Voigt kernel build when it runs alone (launch_chunk_slices)
  int k2s_stages = 5;                // DLA_B200_K2S_STAGES=3/5: ring depth of the contraction when it runs alone
  int overlap_chunks = 16;           // DLA_B200_OVERLAP_CHUNKS: chunks an overlapped batch is split into (at most)
  bool overlap = false;               // DLA_B200_OVERLAP=0: contraction of chunk i NOT run under the Voigt kernel of chunk i + 1
  std::vector<SliceQuasar> h_sq;
  DevBuf d_sq, d_isc, d_csc;
  // slice-path workspace of one chunk: sample planes, table planes and scales, partial sums; two sets when chunks overlap
  struct SliceSet {
    DevBuf W, U, btab, tau, colscale, aux;
    cudaEvent_t planes_ready = nullptr, done = nullptr;
    bool busy = false;   // `done` has been recorded and not yet waited for by the host
  } sset[2];
  cudaStream_t stream2 = nullptr;    // contraction stream of the overlapped slice path (higher priority)
  DevBuf d_trace;                    // DLA_B200_TRACE=<file>: CTA residency trace of the slice kernels (diagnosis)
  std::string trace_path;
  int slice_chunks = 0;
  int num_sms = 148;
  // staged batch
  int64_t Q = 0;
  bool staged = false, ran = false;
  std::vector<QuasarDesc> h_desc;
  std::vector<Chunk> chunks;
  DevBuf d_desc, d_padded, d_pidx, d_flux, d_mu, d_om, d_nv, d_M, d_pix4;
  DevBuf d_sample_ll, d_ll_no, d_ll_dla, d_lp_no, d_lp_dla, d_post, d_map;
  // chunk workspace
  DevBuf d_W, d_U, d_bmat, d_aux, d_ll_chunk, d_G, d_krank;
  dla_timing timing{};
};

#define CU_TRY(ctx, call)                                                                          \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      char b_[512];                                                                                \
      snprintf(b_, sizeof b_, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__,    \
               __LINE__);                                                                          \
      (ctx)->err = b_;                                                                             \
      return e_ == cudaErrorMemoryAllocation ? DLA_ERR_OOM : DLA_ERR_CUDA;                         \
    }                                                                                              \
  } while (0)

#define BAD_ARG(ctx, msg)       \
  do {                          \
    (ctx)->err = (msg);         \
    return DLA_ERR_BAD_ARG;     \
  } while (0)

static inline int64_t round_up(int64_t a, int64_t b) { return (a + b - 1) / b * b; }

// ------------------------------------------------------------------------------------------
// chunk execution
// ------------------------------------------------------------------------------------------
// Column geometry of the contraction: the fused path (k <= 20) has NCOL = 240 columns in one CTA
// tile; the generic-rank path splits roundup(T,8) + roundup(k,8) columns into column blocks.
struct ColGeom {
  bool generic;
  int ncol, voff, nblocks, tpw;
};
static ColGeom col_geom(int k) {
  if (k <= dla::KMAX) return {false, dla::NCOL, dla::VOFF, 1, dla::NT / 2};
  const int T8 = (int)round_up((int64_t)k * (k + 1) / 2, 8);
  const int ntile = (T8 + (int)round_up(k, 8)) / 8;
  const int nblocks = (ntile + dla::NT - 1) / dla::NT;
  const int tpw = std::max(8, (ntile + 2 * nblocks - 1) / (2 * nblocks));
  return {true, nblocks * 2 * tpw * 8 /* zero columns pad the last block */, T8, nblocks, tpw};
}

// contract_block_kernel<TPW> for the runtime tiles-per-warp of a column geometry
template <typename... Args>
static void launch_contract_block(int tpw, dim3 grid, size_t smem, cudaStream_t st, Args... args) {
  switch (tpw) {
#define CB_CASE(T) case T: dla::contract_block_kernel<T><<<grid, dla::K2_THREADS, smem, st>>>(args...); break;
    CB_CASE(8) CB_CASE(9) CB_CASE(10) CB_CASE(11) CB_CASE(12) CB_CASE(13) CB_CASE(14) CB_CASE(15)
#undef CB_CASE
    default: break;
  }
}

static int k1_warps_for(int raw_stride) {
  const size_t budget = 200 * 1024 - 256;
  int w = (int)(budget / ((size_t)raw_stride * 8));
  if (w > dla::K1_THREADS / 32) w = dla::K1_THREADS / 32;
  return w;
}

// timing events per chunk: [0] start, [1] tables done, [2] Voigt done, [3] contraction start, [4] contraction done,
// [5] Cholesky done, [6] reduction done ([0..2] on the compute stream, [3..6] on the stream the contraction runs on)
constexpr int NTEV = 7;

template <typename... Args>
static void launch_k2s(int mode, int stages, unsigned grid, cudaStream_t st, Args... args) {
  namespace sl = dla::sl;
  if (mode == 1 && stages == 3) sl::contract_slices_tiles_kernel<1, 3><<<grid, sl::K2S_THREADS, sl::k2s_smem(3), st>>>(args...);
  else if (mode == 2 && stages == 3) sl::contract_slices_tiles_kernel<2, 3><<<grid, sl::K2S_THREADS, sl::k2s_smem(3), st>>>(args...);
  else if (mode == 1) sl::contract_slices_tiles_kernel<1, 5><<<grid, sl::K2S_THREADS, sl::k2s_smem(5), st>>>(args...);
  else if (mode == 2) sl::contract_slices_tiles_kernel<2, 5><<<grid, sl::K2S_THREADS, sl::k2s_smem(5), st>>>(args...);
  else if (mode == 3) sl::contract_slices_tiles_kernel<3, 5><<<grid, sl::K2S_THREADS, sl::k2s_smem(5), st>>>(args...);
  else if (stages == 3) sl::contract_slices_tiles_kernel<0, 3><<<grid, sl::K2S_THREADS, sl::k2s_smem(3), st>>>(args...);
  else sl::contract_slices_tiles_kernel<0, 5><<<grid, sl::K2S_THREADS, sl::k2s_smem(5), st>>>(args...);
}

// The exact-slice path of one chunk (slices_path.cuh): K0s tables and K1s Voigt -> int8 planes on the compute stream,
// then K2s tcgen05 contraction, K2c Cholesky rows and K3 reduction on the contraction stream.  With overlap on, the
// contraction stream is a second (higher-priority) stream and the plane / table / partial-sum buffers alternate
// between two sets, so the tensor-pipe work of chunk i runs under the FP64-pipe work of chunk i + 1 on the same SMs.
static int launch_chunk_slices(dla_ctx* c, const Inputs& in, const Chunk& ch, int64_t S, int k, const double* d_offset,
                               const double* d_nhi, const int32_t* d_perm, double* sample_ll_chunk, double* ll_no,
                               double* ll_dla, double* lp_no, double* lp_dla, double* post, int32_t* map_index,
                               cudaEvent_t* tev) {
  namespace sl = dla::sl;
  const bool time_it = tev != nullptr;
  const int nsb = (int)((S + 1 + sl::SB - 1) / sl::SB);   // 128-sample blocks
  const sl::SliceLayout L = sl::slice_layout(k);
  const bool ov = c->overlap;
  cudaStream_t sa = c->stream, sb = ov ? c->stream2 : c->stream;
  dla_ctx::SliceSet& set = c->sset[ov ? (c->slice_chunks & 1) : 0];
  c->slice_chunks++;
  int8_t* Aw = set.W.as<int8_t>();
  int8_t* Au = set.U.as<int8_t>();
  int8_t* Bt = set.btab.as<int8_t>();
  double* G = c->d_G.as<double>();
  double* aux = set.aux.as<double>();
  if (ov && set.busy) CU_TRY(c, cudaStreamWaitEvent(sa, set.done, 0));   // the chunk that last used this set has finished
  if (time_it) CU_TRY(c, cudaEventRecord(tev[0], sa));
  {
    NvtxRange r("dla K0s tables");
    dim3 grid((unsigned)((ch.max_npad + 255) / 256), (unsigned)ch.nq);
    dla::pack_pixels_kernel<<<grid, 256, 0, sa>>>(in.desc, ch.q0, in.flux, in.mu, in.om, in.nv, in.pix4);
    sl::pixel_scales_kernel<<<grid, 256, 0, sa>>>(in.desc, ch.q0, in.flux, in.mu, in.om, in.nv, c->d_isc.as<double2>(),
                                                  c->d_csc.as<double2>());
    sl::table_colscale_kernel<<<dim3((unsigned)L.gcols, (unsigned)ch.nq), 128, 0, sa>>>(
        in.desc, ch.q0, in.M, L, c->d_csc.as<double2>(), set.tau.as<double>(), set.colscale.as<double>());
    sl::table_slices_kernel<<<dim3((unsigned)((2 * ch.max_nks * L.gcols + 255) / 256), (unsigned)ch.nq), 256, 0, sa>>>(
        in.desc, in.sq, ch.q0, in.M, L, c->d_csc.as<double2>(), set.tau.as<double>(), Bt);
  }
  if (time_it) CU_TRY(c, cudaEventRecord(tev[1], sa));
  const int nch = 2 * ch.max_nks;
  const int nseg = std::min(dla::K1S_NSEG_MAX, std::max(dla::K1S_NSEG_MIN, (nch + 19) / 20));
  {
    NvtxRange r("dla K1s voigt -> int8 planes");
    const int cmax = (nch + nseg - 1) / nseg;   // 16-pixel chunks per segment
    const int lam_cap = 16 * cmax + 8, px_cap = 16 * cmax;
    const int variant = ov ? 2 : c->k1q_variant;   // 0: 3 CTAs per SM (162 registers), records in shared memory; 1 (default): 4 CTAs (126 registers, no spills), same; 2: 4 CTAs, records from L1
    const size_t smem = (size_t)lam_cap * 8 + (variant == 2 ? 0 : (size_t)px_cap * 48) + (size_t)4 * 3 * sl::K1Q_THREADS * 16;
    if (smem > 100 * 1024) BAD_ARG(c, "a quasar has too many pixels for the Voigt kernel's shared-memory staging");
    const unsigned grid = (unsigned)((int64_t)ch.nq * nsb * nseg);
#define K1Q_ARGS in.desc, in.sq, ch.q0, nsb, S, d_offset, d_nhi, in.padded, in.pix4, c->d_isc.as<double2>(), \
                 c->d_etab256.as<double>(), Aw, Au, aux, lam_cap, px_cap, nseg
    if (variant == 2) sl::voigt_slices_lane_kernel<4, false><<<grid, sl::K1Q_THREADS, smem, sa>>>(K1Q_ARGS);
    else if (variant == 1) sl::voigt_slices_lane_kernel<4, true><<<grid, sl::K1Q_THREADS, smem, sa>>>(K1Q_ARGS);
    else sl::voigt_slices_lane_kernel<3, true><<<grid, sl::K1Q_THREADS, smem, sa>>>(K1Q_ARGS);
#undef K1Q_ARGS
  }
  if (time_it) CU_TRY(c, cudaEventRecord(tev[2], sa));
  if (ov) {
    CU_TRY(c, cudaEventRecord(set.planes_ready, sa));
    CU_TRY(c, cudaStreamWaitEvent(sb, set.planes_ready, 0));
  }
  if (time_it) CU_TRY(c, cudaEventRecord(tev[3], sb));
  {
    NvtxRange r("dla K2s tcgen05 contraction");
    const int64_t ntiles = (int64_t)ch.nq * nsb * L.nunits;
    if (ntiles > (int64_t)INT32_MAX) BAD_ARG(c, "too many contraction tiles in one chunk");
    const unsigned grid = (unsigned)std::min<int64_t>(ntiles, c->num_sms);
    launch_k2s(c->k2s_mode, ov ? 3 : c->k2s_stages, grid, sb, in.sq, ch.q0, ch.nq, nsb, L, (const int8_t*)Aw, (const int8_t*)Au,
               (const int8_t*)Bt, (const double*)set.colscale.as<double>(), G);
  }
  if (time_it) CU_TRY(c, cudaEventRecord(tev[4], sb));
  {
    NvtxRange r("dla K2c cholesky + K3 reduce");
    if (k <= dla::KMAX) {
      dim3 grid((unsigned)(nsb * (sl::SB / sl::K2C_ROWS)), (unsigned)ch.nq);
      sl::chol_rows_kernel<<<grid, sl::K2C_ROWS * 4, sl::K2C_SMEM, sb>>>(in.desc, ch.q0, nsb, S, G, aux, d_perm, sample_ll_chunk,
                                                                        ll_no, k, nseg);
    } else {
      // rows of G are [S_pad128][gcols]: the generic bordered Cholesky reads them with its 64-sample block count doubled
      dim3 g2((unsigned)((S + 1 + dla::CHG_WARPS - 1) / dla::CHG_WARPS), (unsigned)ch.nq);
      dla::chol_generic_kernel<<<g2, dla::CHG_WARPS * 32, dla::chg_smem_bytes(k), sb>>>(
          in.desc, ch.q0, nsb * (sl::SB / dla::TS), S, G, aux, d_perm, sample_ll_chunk, ll_no, k, L.gcols, L.voff, nseg);
    }
  }
  if (time_it) CU_TRY(c, cudaEventRecord(tev[5], sb));
  if (S > 0 && ll_dla) {
    dla::reduce_kernel<<<(unsigned)ch.nq, 256, 0, sb>>>(in.desc, ch.q0, S, sample_ll_chunk, ll_no, ll_dla, lp_no,
                                                         lp_dla, post, map_index);
  }
  if (time_it) CU_TRY(c, cudaEventRecord(tev[6], sb));
  if (ov) {
    CU_TRY(c, cudaEventRecord(set.done, sb));
    set.busy = true;
  }
  CU_TRY(c, cudaGetLastError());
  c->timing.launches += (S > 0 && ll_dla) ? 8 : 7;
  c->timing.contraction = 1;
  return DLA_OK;
}

// Runs K0..K3 for one chunk of quasars on ctx->stream. sample_ll_chunk: [nq][S].
static int launch_chunk(dla_ctx* c, const Inputs& in, const Chunk& ch, int64_t S, int k, int num_lines,
                        const double* d_offset, const double* d_nhi, const int32_t* d_perm, double* sample_ll_chunk, double* ll_no,
                        double* ll_dla, double* lp_no, double* lp_dla, double* post, int32_t* map_index,
                        cudaEvent_t* tev /* NTEV timing events or nullptr */) {
  const bool time_it = tev != nullptr;
  if (ch.slices) return launch_chunk_slices(c, in, ch, S, k, d_offset, d_nhi, d_perm, sample_ll_chunk, ll_no, ll_dla, lp_no,
                                            lp_dla, post, map_index, tev);
  const int nsb = (int)((S + 1 + dla::TS - 1) / dla::TS);
  const ColGeom gen = col_geom(k);
  double* W = c->d_W.as<double>();
  double* U = c->d_U.as<double>();
  double* B = c->d_bmat.as<double>();
  double* aux = c->d_aux.as<double>();
  cudaStream_t st = c->stream;
  for (auto& set : c->sset)   // overlapped slice chunks still in flight share the per-chunk output buffers
    if (set.busy) {
      CU_TRY(c, cudaStreamWaitEvent(st, set.done, 0));
      set.busy = false;
    }
  if (time_it) CU_TRY(c, cudaEventRecord(tev[0], st));
  {
    dim3 grid((unsigned)(((int64_t)(ch.max_npad / 4) * gen.ncol + 255) / 256), (unsigned)ch.nq);
    dla::build_bmat_kernel<<<grid, 256, 0, st>>>(in.desc, ch.q0, in.M, B, k, gen.ncol, gen.voff);
  }
  {
    dim3 grid((unsigned)((ch.max_npad + 255) / 256), (unsigned)ch.nq);
    dla::pack_pixels_kernel<<<grid, 256, 0, st>>>(in.desc, ch.q0, in.flux, in.mu, in.om, in.nv, in.pix4);
  }
  if (time_it) CU_TRY(c, cudaEventRecord(tev[1], st));
  int nseg = 1;  // partial (q0, log|D|) sums per sample that K1 leaves in aux
  const bool fused = c->use_fused && ch.all_ident && num_lines == 3 && !gen.generic;
  if (fused) {
    // K1 runs inside the contraction kernel (producer warps)
  } else if (ch.all_ident && num_lines == 3 && !c->force_pixel_lane_k1) {
    nseg = std::min(dla::K1S_NSEG_MAX, std::max(dla::K1S_NSEG_MIN, (ch.max_npad / 4 + 79) / 80));
    const int gmax = (ch.max_npad / 4 + nseg - 1) / nseg + 1;  // groups of 4 pixels per segment
    const int lam_cap = 4 * gmax + 8;
    const size_t smem = ((size_t)lam_cap + 16 * (size_t)gmax) * 8;
    if (smem > 100 * 1024) BAD_ARG(c, "a quasar has too many pixels for the Voigt kernel's shared-memory staging");
    dla::voigt_weights_lane_kernel<<<(unsigned)((int64_t)ch.nq * nsb * nseg), dla::K1S_THREADS, smem, st>>>(
        in.desc, ch.q0, nsb, S, d_offset, d_nhi, in.padded, in.pix4, c->d_etab256.as<double>(), W, U, aux, lam_cap, nseg);
  } else if (num_lines == 3 && !c->force_pixel_lane_k1) {
    // masked gather (profile_index[j] != j for some quasar of the chunk): lane kernel over profile positions
    {
      dim3 grid((unsigned)((ch.max_P - 5 + 255) / 256), (unsigned)ch.nq);
      dla::kept_rank_kernel<<<grid, 256, 0, st>>>(in.desc, ch.q0, in.pidx, c->d_krank.as<int32_t>());
    }
    const int ngrp = (ch.max_P - 6 + 3) / 4;
    nseg = std::min(dla::K1S_NSEG_MAX, std::max(dla::K1S_NSEG_MIN, (ngrp + 79) / 80));
    const int gmax = (ngrp + nseg - 1) / nseg + 1;
    const int lam_cap = 4 * gmax + 8, px_cap = 4 * gmax;
    const size_t smem = ((size_t)lam_cap + 4 * (size_t)px_cap) * 8 + ((size_t)4 * gmax + 2) * 4;
    if (smem > 100 * 1024) BAD_ARG(c, "a quasar has too many pixels for the Voigt kernel's shared-memory staging");
    dla::voigt_weights_lane_gather_kernel<<<(unsigned)((int64_t)ch.nq * nsb * nseg), dla::K1S_THREADS, smem, st>>>(
        in.desc, ch.q0, nsb, S, d_offset, d_nhi, in.padded, in.pix4, c->d_krank.as<int32_t>(),
        c->d_etab256.as<double>(), W, U, aux, lam_cap, px_cap, nseg);
    c->timing.launches += 1;
  } else {
    const int raw_stride = (int)round_up(ch.max_P, 2) + 2;  // +2: the 16-byte window loads of the tail may touch one double past P
    const int warps = k1_warps_for(raw_stride);
    if (warps < 1) BAD_ARG(c, "a quasar has too many padded pixels for the Voigt kernel's shared-memory row");
    const size_t smem = (size_t)warps * raw_stride * 8 + 32 * 8;
    dla::voigt_weights_kernel<<<(unsigned)((int64_t)ch.nq * nsb), warps * 32, smem, st>>>(
        in.desc, ch.q0, nsb, S, d_offset, d_nhi, in.padded, in.pidx, in.pix4, W, U, aux,
        num_lines, raw_stride);
  }
  if (time_it) CU_TRY(c, cudaEventRecord(tev[2], st));
  if (time_it) CU_TRY(c, cudaEventRecord(tev[3], st));
  if (fused) {
    const int64_t ntiles = (int64_t)ch.nq * nsb;
    const unsigned grid = (unsigned)std::min<int64_t>(ntiles, c->num_sms);
    dla::fused_voigt_contract_kernel<<<grid, dla::KF_THREADS, dla::KF_SMEM_BYTES, st>>>(
        in.desc, ch.q0, nsb, (int)ntiles, S, d_offset, d_nhi, in.padded, in.pix4, c->d_etab256.as<double>(), B, d_perm,
        sample_ll_chunk, ll_no, k);
  } else if (gen.generic) {
    const size_t smem = (size_t)dla::K2_STAGES * dla::STAGE_DOUBLES * 8 + 64;
    dim3 grid((unsigned)((int64_t)ch.nq * nsb), (unsigned)gen.nblocks);
    launch_contract_block(gen.tpw, grid, smem, st, in.desc, ch.q0, nsb, (const double*)W, (const double*)U,
                          (const double*)B, c->d_G.as<double>(), gen.ncol, gen.voff / 8);
    dim3 g2((unsigned)((S + 1 + dla::CHG_WARPS - 1) / dla::CHG_WARPS), (unsigned)ch.nq);
    const size_t sm2 = dla::chg_smem_bytes(k);
    dla::chol_generic_kernel<<<g2, dla::CHG_WARPS * 32, sm2, st>>>(in.desc, ch.q0, nsb, S, c->d_G.as<double>(), aux,
                                                                    d_perm, sample_ll_chunk, ll_no, k, gen.ncol, gen.voff, nseg);
  } else if (!c->force_tile_k2) {
    const int64_t ntiles = (int64_t)ch.nq * nsb;
    const unsigned grid = (unsigned)std::min<int64_t>(ntiles, c->num_sms);
    if (c->k2_rowsplit)
      dla::contract_chol_rowsplit_kernel<<<grid, dla::K2P_THREADS, dla::K2P_SMEM_BYTES, st>>>(
          in.desc, ch.q0, nsb, (int)ntiles, S, W, U, B, aux, d_perm, sample_ll_chunk, ll_no, k, nseg);
    else
      dla::contract_chol_persistent_kernel<<<grid, dla::K2P_THREADS, dla::K2P_SMEM_BYTES, st>>>(
          in.desc, ch.q0, nsb, (int)ntiles, S, W, U, B, aux, d_perm, sample_ll_chunk, ll_no, k, nseg);
  } else {
    const size_t smem = (size_t)dla::K2_STAGES * dla::STAGE_DOUBLES * 8 + 64;
    dla::contract_chol_kernel<<<(unsigned)((int64_t)ch.nq * nsb), dla::K2_THREADS, smem, st>>>(
        in.desc, ch.q0, nsb, S, W, U, B, aux, d_perm, sample_ll_chunk, ll_no, k, nseg);
  }
  if (time_it) CU_TRY(c, cudaEventRecord(tev[4], st));
  if (time_it) CU_TRY(c, cudaEventRecord(tev[5], st));  // the Cholesky is part of the contraction kernels on this path
  if (S > 0 && ll_dla) {
    dla::reduce_kernel<<<(unsigned)ch.nq, 256, 0, st>>>(in.desc, ch.q0, S, sample_ll_chunk, ll_no, ll_dla, lp_no,
                                                         lp_dla, post, map_index);
  }
  if (time_it) CU_TRY(c, cudaEventRecord(tev[6], st));
  CU_TRY(c, cudaGetLastError());
  c->timing.launches += ((S > 0 && ll_dla) ? 5 : 4) + (gen.generic ? 1 : 0);
  return DLA_OK;
}

// after the stream has been synchronised: per-kernel device times summed over the chunks
static int collect_timing(dla_ctx* c, size_t nchunks) {
  for (size_t i = 0; i < nchunks; ++i) {
    cudaEvent_t* e = &c->ev_time[NTEV * i];
    float ms;
    CU_TRY(c, cudaEventElapsedTime(&ms, e[0], e[1])); c->timing.ms_phi += ms;
    CU_TRY(c, cudaEventElapsedTime(&ms, e[1], e[2])); c->timing.ms_voigt += ms;
    CU_TRY(c, cudaEventElapsedTime(&ms, e[3], e[4])); c->timing.ms_contract += ms;
    CU_TRY(c, cudaEventElapsedTime(&ms, e[4], e[5])); c->timing.ms_chol += ms;
    CU_TRY(c, cudaEventElapsedTime(&ms, e[5], e[6])); c->timing.ms_reduce += ms;
    // chunks may overlap in time (contraction of chunk i under the Voigt kernel of chunk i + 1): the total is the
    // span from the first chunk's start to the last completion, not a sum
    CU_TRY(c, cudaEventElapsedTime(&ms, c->ev_time[0], e[6]));
    c->timing.ms_total = std::max(c->timing.ms_total, (double)ms);
  }
  return DLA_OK;
}

// both compute streams idle; the slice sets are free again
static int sync_compute(dla_ctx* c) {
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream2));
  for (auto& set : c->sset) set.busy = false;
  c->slice_chunks = 0;
  return DLA_OK;
}

static int ensure_events(dla_ctx* c, size_t nchunks) {
  while (c->ev_time.size() < NTEV * nchunks) {
    cudaEvent_t e;
    CU_TRY(c, cudaEventCreate(&e));
    c->ev_time.push_back(e);
  }
  while (c->ev_sync.size() < 2 * nchunks) {
    cudaEvent_t e;
    CU_TRY(c, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    c->ev_sync.push_back(e);
  }
  return DLA_OK;
}

static int ensure_workspace(dla_ctx* c, int k, int64_t wu_doubles, int64_t bmat_doubles, int nq, int64_t S, bool need_ll_chunk,
                            int64_t a_bytes = 0, int64_t b_bytes = 0, int nq_slices = 0) {
  const int64_t S_pad = round_up(S + 1, dla::TS), S_pad128 = round_up(S + 1, dla::sl::SB);
  CU_TRY(c, c->d_W.ensure((size_t)wu_doubles * 8));
  CU_TRY(c, c->d_U.ensure((size_t)wu_doubles * 8));
  const ColGeom gen = col_geom(k);
  CU_TRY(c, c->d_bmat.ensure((size_t)(bmat_doubles / dla::NCOL + 1) * gen.ncol * 8));
  CU_TRY(c, c->d_aux.ensure((size_t)nq * 2 * dla::K1S_NSEG_MAX * S_pad128 * 8));
  if (nq_slices > 0) {
    const size_t gcols = (size_t)dla::sl::slice_layout(k).gcols;
    CU_TRY(c, c->d_G.ensure((size_t)nq_slices * S_pad128 * gcols * 8));
    for (int i = 0; i < (c->overlap ? 2 : 1); ++i) {
      dla_ctx::SliceSet& set = c->sset[i];
      CU_TRY(c, set.W.ensure((size_t)std::max<int64_t>(a_bytes, 32)));
      CU_TRY(c, set.U.ensure((size_t)std::max<int64_t>(a_bytes, 32)));
      CU_TRY(c, set.btab.ensure((size_t)std::max<int64_t>(b_bytes, 32)));
      CU_TRY(c, set.tau.ensure((size_t)nq_slices * gcols * 8));
      CU_TRY(c, set.colscale.ensure((size_t)nq_slices * gcols * 8));
      CU_TRY(c, set.aux.ensure((size_t)nq_slices * 2 * dla::K1S_NSEG_MAX * S_pad128 * 8));
    }
  }
  if (gen.generic) CU_TRY(c, c->d_G.ensure((size_t)nq * S_pad * gen.ncol * 8));
  if (need_ll_chunk) CU_TRY(c, c->d_ll_chunk.ensure((size_t)std::max<int64_t>(1, (int64_t)nq * S) * 8));
  return DLA_OK;
}

// ------------------------------------------------------------------------------------------
// ABI
// ------------------------------------------------------------------------------------------
// ------------------------------------------------------------------------------------------
// dla_voigt_noalloc: scratch that is never allocated.  One file-scope device array per GPU that loads the library
// (384 KB, held until the process exits): [0, MAX) staged wavelengths, [MAX, 2 MAX) raw transmissions (the kernel's
// scratch, voigt.c:282-292), [2 MAX, 3 MAX) the broadened profile.  In host code the array's NAME is only a stand-in;
// its device address comes from cudaGetSymbolAddress, once per device, with that device current (dla_create).
// ------------------------------------------------------------------------------------------
static __device__ __align__(256) double g_noalloc_scratch[3 * DLA_NOALLOC_MAX_POINTS];
static std::mutex g_noalloc_mu;
static double* g_noalloc_addr[64] = {};
// `device` must be the current device
static cudaError_t noalloc_scratch_address(int device, double** out) {
  if (device < 0 || device >= 64) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_addr[device]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    g_noalloc_addr[device] = static_cast<double*>(p);
  }
  *out = g_noalloc_addr[device];
  return cudaSuccess;
}

extern "C" {

int dla_abi_version(void) { return DLA_B200_ABI_VERSION; }

int dla_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

const char* dla_last_error(const dla_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int dla_create(dla_ctx** out, int device, const dla_params* p) {
  if (!out || !p) {
    g_create_error = "dla_create: NULL argument";
    return DLA_ERR_BAD_ARG;
  }
  *out = nullptr;
  if (p->k < 1 || p->k > dla::KGEN_MAX) {
    g_create_error = "dla_create: k must be in 1..64 (1..20 fused epilogue, 21..64 column-block path)";
    return DLA_ERR_BAD_ARG;
  }
  if (p->num_lines < 1 || p->num_lines > DLA_MAX_LINES) {
    g_create_error = "dla_create: num_lines must be in 1..31 (voigt.c:16)";
    return DLA_ERR_BAD_ARG;
  }
  if (p->num_samples < 0) {
    g_create_error = "dla_create: num_samples < 0";
    return DLA_ERR_BAD_ARG;
  }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    g_create_error = std::string("dla_create: no CUDA device available (") +
                     (e != cudaSuccess ? cudaGetErrorString(e) : "device count 0") +
                     "); this library has no CPU fallback";
    return DLA_ERR_NO_DEVICE;
  }
  if (device < 0 || device >= ndev) {
    g_create_error = "dla_create: device index out of range";
    return DLA_ERR_BAD_ARG;
  }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess || prop.major != 10) {
    g_create_error = "dla_create: device is not sm_100 (kernels are built for sm_100a only)";
    return DLA_ERR_NO_DEVICE;
  }
  dla_ctx* c = new dla_ctx();
  c->device = device;
  c->p = *p;
  if (c->p.workspace_bytes <= 0) c->p.workspace_bytes = (int64_t)16 << 30;
#define CREATE_TRY(call)                                                                   \
  do {                                                                                     \
    cudaError_t e2_ = (call);                                                              \
    if (e2_ != cudaSuccess) {                                                              \
      g_create_error = std::string("dla_create: " #call " failed: ") + cudaGetErrorString(e2_); \
      delete c;                                                                            \
      return DLA_ERR_CUDA;                                                                 \
    }                                                                                      \
  } while (0)
  CREATE_TRY(cudaSetDevice(device));
  CREATE_TRY(noalloc_scratch_address(device, &c->noalloc_scratch));
  CREATE_TRY(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  CREATE_TRY(cudaStreamCreateWithFlags(&c->s_in, cudaStreamNonBlocking));
  CREATE_TRY(cudaStreamCreateWithFlags(&c->s_out, cudaStreamNonBlocking));
  {
    int prio_lo = 0, prio_hi = 0;
    CREATE_TRY(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    // the persistent contraction must get its one CTA per SM as soon as a Voigt CTA of the next chunk retires
    CREATE_TRY(cudaStreamCreateWithPriority(&c->stream2, cudaStreamNonBlocking, prio_hi));
    for (auto& set : c->sset) {
      CREATE_TRY(cudaEventCreateWithFlags(&set.planes_ready, cudaEventDisableTiming));
      CREATE_TRY(cudaEventCreateWithFlags(&set.done, cudaEventDisableTiming));
    }
  }
  for (auto& ev : c->ev) CREATE_TRY(cudaEventCreate(&ev));
  // per-line constants, computed as libcerf's voigt() does: (x + i gamma)/sqrt(2)/sigma, /sqrt(2 pi)/sigma
  {
    double gam[DLA_MAX_LINES], lc[DLA_MAX_LINES];
    CREATE_TRY(cudaMemcpyFromSymbol(gam, dla_gammas, sizeof gam));
    CREATE_TRY(cudaMemcpyFromSymbol(lc, dla_leading_constants, sizeof lc));
    dla::LineConsts h;
    memset(&h, 0, sizeof h);
    const double sigma = DLA_SIGMA, s2pi = 2.50662827463100050242;
    const double T = 2.0 * sigma * sigma;
    for (int j = 0; j < DLA_MAX_LINES; ++j) {
      h.y[j] = gam[j] / sqrt(2.0) / sigma;
      h.coef[j] = lc[j] / s2pi / sigma;
      h.cwT[j] = h.coef[j] * h.y[j] * DLA_INV_SQRTPI * T;
      h.y2T[j] = h.y[j] * h.y[j] * T;
      if (!(h.y[j] > DLA_FADD_YMIN && h.y[j] < DLA_FADD_YMAX)) {
        g_create_error = "dla_create: Lorentz width outside the validated Faddeeva domain";
        delete c;
        return DLA_ERR_BAD_ARG;
      }
    }
    {
      const double g[7] = {1.0, 1.5, 3.75, 13.125, 59.0625, 324.84375, 2111.484375};  // (2n+1)!!/2^n
      double tp = 1.0;
      for (int n = 0; n < 7; ++n) { h.gq[n] = g[n] * tp; tp *= T; }
    }
    h.fiveT = 5.0 * T;
    h.vnear2 = (DLA_FADD_XNEAR * sqrt(2.0) * sigma) * (DLA_FADD_XNEAR * sqrt(2.0) * sigma);
    h.inv_sqrt2_sigma = 1.0 / (sqrt(2.0) * sigma);
    h.c_cgs = DLA_C_CGS;
    h.k256 = 256.0 * 1.4426950408889634074;
    h.nln2hi256 = -6.93147180369123816490e-01 / 256.0;
    h.nln2lo256 = -1.90821492927058770002e-10 / 256.0;
    h.nln2_256 = -0.693147180559945309417232 / 256.0;
    for (int j = 0; j < DLA_MAX_LINES; ++j) h.g1p[j] = h.gq[1] - h.y2T[j];
    h.k32 = 32.0 * 1.4426950408889634074;  // 32 / ln 2
    h.magic = 6755399441055744.0;           // 1.5 * 2^52: adding it rounds to the nearest integer
    h.nln2hi32 = -6.93147180369123816490e-01 / 32.0;  // exact scalings of the hi/lo split of ln 2
    h.nln2lo32 = -1.90821492927058770002e-10 / 32.0;
    {
      double f = 1.0;
      for (int n = 0; n < 7; ++n) { if (n > 0) f *= n; h.ec[n] = 1.0 / f; }
      for (int i = 0; i < 32; ++i) h.etab[i] = exp2(i / 32.0);
    }
    {
      auto hi_of = [&](double xthr) {
        const double v = xthr * sqrt(2.0) * sigma;
        const double v2 = v * v;
        long long bits;
        memcpy(&bits, &v2, 8);
        return (int)(bits >> 32) + 1;  // strictly above the threshold: the series is only used beyond it
      };
      h.hi_near = hi_of(DLA_FADD_XNEAR);
      h.hi_64 = hi_of(64.0);
      h.hi_256 = hi_of(256.0);
      h.hi_64x = hi_of(64.5);
      {  // high word of 1 / v^2 at 260 widths: a pair with hi(1/v^2) below it is safely beyond 256
        const double v = 260.0 * sqrt(2.0) * sigma, r = 1.0 / (v * v);
        long long bits;
        memcpy(&bits, &r, 8);
        h.hi_sp256 = (int)(bits >> 32);
      }
    }
    CREATE_TRY(cudaMemcpyToSymbol(dla::g_lines, &h, sizeof h));
    double e256[256];
    for (int i = 0; i < 256; ++i) e256[i] = exp2(i / 256.0);
    CREATE_TRY(c->d_etab256.ensure(sizeof e256));
    CREATE_TRY(cudaMemcpy(c->d_etab256.p, e256, sizeof e256, cudaMemcpyHostToDevice));
    const char* fu = getenv("DLA_B200_FUSED");
    c->use_fused = fu && strcmp(fu, "1") == 0;
    const char* k2 = getenv("DLA_B200_K2");
    c->force_tile_k2 = k2 && strcmp(k2, "tile") == 0;
    c->k2_rowsplit = k2 && strcmp(k2, "rowsplit") == 0;
    c->num_sms = prop.multiProcessorCount;
    const char* ct = getenv("DLA_B200_CONTRACT");
    c->use_slices = !(ct && strcmp(ct, "dmma") == 0) && !(k2 && strcmp(k2, "dmma") == 0) && !c->use_fused &&
                    !c->force_tile_k2 && !c->k2_rowsplit;
    const char* km = getenv("DLA_B200_K2S_MODE");
    c->k2s_mode = km ? atoi(km) : 0;
    const char* kq = getenv("DLA_B200_K1Q");
    if (kq) c->k1q_variant = std::max(0, std::min(2, atoi(kq)));
    const char* ks = getenv("DLA_B200_K2S_STAGES");
    c->k2s_stages = (ks && atoi(ks) == 3) ? 3 : 5;
    const char* ov = getenv("DLA_B200_OVERLAP");
    c->overlap = ov && strcmp(ov, "1") == 0;   // measured: no gain (the two kernels contend for shared-memory bandwidth)
    const char* tp = getenv("DLA_B200_TRACE");
    if (tp && *tp) {
      c->trace_path = tp;
      CREATE_TRY(c->d_trace.ensure((size_t)dla::sl::TRACE_CAP * 32));
      unsigned long long* tptr = c->d_trace.as<unsigned long long>();
      CREATE_TRY(cudaMemcpyToSymbol(dla::sl::g_trace, &tptr, sizeof tptr));
    }
    const char* oc = getenv("DLA_B200_OVERLAP_CHUNKS");
    if (oc && atoi(oc) > 0) c->overlap_chunks = atoi(oc);
    const char* k1 = getenv("DLA_B200_K1");
    c->force_pixel_lane_k1 = k1 && strcmp(k1, "pixel") == 0;
    if (c->force_pixel_lane_k1) c->use_slices = false;
  }
  CREATE_TRY(cudaFuncSetAttribute(dla::sl::contract_slices_tiles_kernel<0, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dla::sl::k2s_smem(5)));
  CREATE_TRY(cudaFuncSetAttribute(dla::sl::contract_slices_tiles_kernel<0, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dla::sl::k2s_smem(3)));
  CREATE_TRY(cudaFuncSetAttribute(dla::sl::contract_slices_tiles_kernel<1, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dla::sl::k2s_smem(5)));
  CREATE_TRY(cudaFuncSetAttribute(dla::sl::contract_slices_tiles_kernel<2, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dla::sl::k2s_smem(5)));
  CREATE_TRY(cudaFuncSetAttribute(dla::sl::contract_slices_tiles_kernel<3, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dla::sl::k2s_smem(5)));
  CREATE_TRY(cudaFuncSetAttribute(dla::sl::contract_slices_tiles_kernel<1, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dla::sl::k2s_smem(3)));
  CREATE_TRY(cudaFuncSetAttribute(dla::sl::contract_slices_tiles_kernel<2, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dla::sl::k2s_smem(3)));
  // the contraction of one chunk and the Voigt kernel of the next share SMs: both ask for the largest shared-memory
  // carveout, otherwise the SM is configured for the first one alone (132 KB for the 3-stage contraction) and the other
  // kernel's CTAs wait for it to drain
  CREATE_TRY(cudaFuncSetAttribute(dla::sl::contract_slices_tiles_kernel<0, 3>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  CREATE_TRY(cudaFuncSetAttribute(dla::sl::contract_slices_tiles_kernel<1, 3>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  CREATE_TRY(cudaFuncSetAttribute(dla::sl::contract_slices_tiles_kernel<2, 3>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  CREATE_TRY(cudaFuncSetAttribute(dla::sl::voigt_slices_lane_kernel<4, false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  CREATE_TRY(cudaFuncSetAttribute(dla::sl::chol_rows_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
  CREATE_TRY(cudaFuncSetAttribute(dla::sl::chol_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dla::sl::K2C_SMEM));
  CREATE_TRY(cudaFuncSetAttribute(dla::sl::voigt_slices_lane_kernel<3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  CREATE_TRY(cudaFuncSetAttribute(dla::sl::voigt_slices_lane_kernel<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  CREATE_TRY(cudaFuncSetAttribute(dla::sl::voigt_slices_lane_kernel<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  CREATE_TRY(cudaFuncSetAttribute(dla::contract_chol_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  dla::K2_STAGES * dla::STAGE_DOUBLES * 8 + 64));
  CREATE_TRY(cudaFuncSetAttribute(dla::fused_voigt_contract_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)dla::KF_SMEM_BYTES));
  CREATE_TRY(cudaFuncSetAttribute(dla::contract_chol_rowsplit_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)dla::K2P_SMEM_BYTES));
  CREATE_TRY(cudaFuncSetAttribute(dla::contract_chol_persistent_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)dla::K2P_SMEM_BYTES));
#define CB_ATTR(T)                                                                                         \
  CREATE_TRY(cudaFuncSetAttribute(dla::contract_block_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                  dla::K2_STAGES * dla::STAGE_DOUBLES * 8 + 64));
  CB_ATTR(8) CB_ATTR(9) CB_ATTR(10) CB_ATTR(11) CB_ATTR(12) CB_ATTR(13) CB_ATTR(14) CB_ATTR(15)
#undef CB_ATTR
  CREATE_TRY(cudaFuncSetAttribute(dla::train_chol_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  dla::TRC_WARPS * ((dla::TRAIN_KMAX + 1) * (dla::TRAIN_KMAX + 2) +
                                                    dla::TRAIN_KMAX * (dla::TRAIN_KMAX + 1) + dla::TRAIN_KMAX) * 8));
  CREATE_TRY(cudaFuncSetAttribute(dla::chol_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  (int)dla::chg_smem_bytes(dla::KGEN_MAX)));
  CREATE_TRY(cudaFuncSetAttribute(dla::voigt_weights_lane_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  CREATE_TRY(cudaFuncSetAttribute(dla::voigt_weights_lane_gather_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
  CREATE_TRY(cudaFuncSetAttribute(dla::voigt_weights_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
#undef CREATE_TRY
  *out = c;
  return DLA_OK;
}

void dla_destroy(dla_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  if (c->stream2) cudaStreamSynchronize(c->stream2);
  for (DevBuf* b : {&c->d_offset, &c->d_nhi, &c->d_perm, &c->d_etab256, &c->d_desc, &c->d_padded, &c->d_pidx, &c->d_flux, &c->d_mu, &c->d_om,
                    &c->d_nv, &c->d_M, &c->d_pix4, &c->d_sample_ll, &c->d_ll_no, &c->d_ll_dla, &c->d_lp_no, &c->d_lp_dla,
                    &c->d_post, &c->d_map, &c->d_W, &c->d_U, &c->d_bmat, &c->d_aux, &c->d_ll_chunk, &c->d_G, &c->d_krank,
                    &c->d_mrest, &c->d_mmu, &c->d_mM, &c->d_mlo, &c->d_lamk, &c->d_zq,
                    &c->d_sq, &c->d_isc, &c->d_csc, &c->d_trace,
                    &c->t_y, &c->t_l, &c->t_nv, &c->t_x, &c->t_om2, &c->t_aux, &c->t_E, &c->t_fq, &c->t_beta, &c->t_P,
                    &c->t_A1, &c->t_S, &c->t_gp, &c->t_g, &c->t_gsc, &c->t_desc})
    b->release();
  for (DevBuf& b : c->seam) b.release();
  for (auto& ev : c->ev)
    if (ev) cudaEventDestroy(ev);
  for (auto& st : c->sstage)
    for (PinBuf* b : {&st.lamk, &st.fk, &st.nk, &st.padded, &st.pidx}) b->release();
  for (auto& set : c->sset) {
    for (DevBuf* b : {&set.W, &set.U, &set.btab, &set.tau, &set.colscale, &set.aux}) b->release();
    if (set.planes_ready) cudaEventDestroy(set.planes_ready);
    if (set.done) cudaEventDestroy(set.done);
  }
  if (c->stream2) cudaStreamDestroy(c->stream2);
  for (auto& ev : c->ev_time) cudaEventDestroy(ev);
  for (auto& ev : c->ev_sync) cudaEventDestroy(ev);
  if (c->s_in) cudaStreamDestroy(c->s_in);
  if (c->s_out) cudaStreamDestroy(c->s_out);
  if (c->stream) cudaStreamDestroy(c->stream);
  delete c;
}

static int dla_set_samples_impl(dla_ctx* c, const double* offset_samples, const double* nhi_samples) {
  if (!c) return DLA_ERR_BAD_ARG;
  if (!offset_samples || !nhi_samples) BAD_ARG(c, "dla_set_samples: NULL pointer");
  CU_TRY(c, cudaSetDevice(c->device));
  const size_t bytes = (size_t)std::max<int64_t>(1, c->p.num_samples) * 8;
  {
    // a caller that passes the same dla_samples.mat arrays with every spectrum (the MEX gateways) pays two memcmp's
    const size_t nb = (size_t)c->p.num_samples * 8;
    if (c->have_samples && c->h_set_offset.size() * 8 == nb && c->h_set_nhi.size() * 8 == nb &&
        (nb == 0 || (memcmp(c->h_set_offset.data(), offset_samples, nb) == 0 && memcmp(c->h_set_nhi.data(), nhi_samples, nb) == 0)))
      return DLA_OK;
    c->have_samples = false;
  }
  CU_TRY(c, c->d_offset.ensure(bytes));
  CU_TRY(c, c->d_nhi.ensure(bytes));
  CU_TRY(c, c->d_perm.ensure((size_t)std::max<int64_t>(1, c->p.num_samples) * 4));
  if (c->p.num_samples > 0) {
    // stable sort by offset (NaN last): the result of a hypothesis does not depend on its position
    // up to the series-length choice it shares with its warp (<= 3e-13 relative either way)
    const int64_t S = c->p.num_samples;
    std::vector<int32_t> perm((size_t)S);
    for (int64_t i = 0; i < S; ++i) perm[(size_t)i] = (int32_t)i;
    std::stable_sort(perm.begin(), perm.end(), [&](int32_t a, int32_t b) {
      const double x = offset_samples[a], y = offset_samples[b];
      if (x != x) return false;
      if (y != y) return true;
      return x < y;
    });
    std::vector<double> so((size_t)S), sn((size_t)S);
    for (int64_t i = 0; i < S; ++i) {
      so[(size_t)i] = offset_samples[perm[(size_t)i]];
      sn[(size_t)i] = nhi_samples[perm[(size_t)i]];
    }
    CU_TRY(c, cudaMemcpyAsync(c->d_offset.p, so.data(), (size_t)S * 8, cudaMemcpyHostToDevice, c->stream));
    CU_TRY(c, cudaMemcpyAsync(c->d_nhi.p, sn.data(), (size_t)S * 8, cudaMemcpyHostToDevice, c->stream));
    CU_TRY(c, cudaMemcpyAsync(c->d_perm.p, perm.data(), (size_t)S * 4, cudaMemcpyHostToDevice, c->stream));
    CU_TRY(c, cudaStreamSynchronize(c->stream));
  }
  c->h_set_offset.assign(offset_samples, offset_samples + c->p.num_samples);
  c->h_set_nhi.assign(nhi_samples, nhi_samples + c->p.num_samples);
  c->have_samples = true;
  return DLA_OK;
}

// Validation, per-quasar descriptors, chunk plan and device buffers for a batch.  No copies yet.
static int plan_batch(dla_ctx* c, int64_t Q, const int64_t* pix_ptr, const int64_t* pad_ptr,
                      const double* padded_wavelengths, const int32_t* profile_index, const double* flux,
                      const double* mu, const double* omega2, const double* noise_variance, const double* M,
                      const double* min_z_dla, const double* max_z_dla, const double* log_prior_dla,
                      const double* log_prior_no_dla, int max_pipe_chunks) {
  c->staged = false;
  c->ran = false;
  if (Q < 0 || Q > (int64_t)INT32_MAX) BAD_ARG(c, "dla_stage_batch: Q out of range");
  if (!pix_ptr || !pad_ptr || !padded_wavelengths || !profile_index || !flux || !mu || !omega2 || !noise_variance ||
      !M || !min_z_dla || !max_z_dla || !log_prior_dla || !log_prior_no_dla)
    BAD_ARG(c, "dla_stage_batch: NULL input pointer");
  if (!c->have_samples) {
    c->err = "dla_stage_batch: dla_set_samples has not been called";
    return DLA_ERR_STATE;
  }
  CU_TRY(c, cudaSetDevice(c->device));
  const int64_t S = c->p.num_samples;
  const int64_t S_pad = round_up(S + 1, dla::TS);
  c->h_desc.assign((size_t)Q, QuasarDesc{});
  for (int64_t q = 0; q < Q; ++q) {
    const int64_t n = pix_ptr[q + 1] - pix_ptr[q], P = pad_ptr[q + 1] - pad_ptr[q];
    if (n < 0 || P < 7 || n > P - 6 || P > (1 << 24)) BAD_ARG(c, "dla_stage_batch: bad CSR (need 0 <= n_q <= P_q - 6)");
    QuasarDesc& d = c->h_desc[(size_t)q];
    d.pix_off = pix_ptr[q];
    d.pad_off = pad_ptr[q];
    d.n = (int32_t)n;
    d.P = (int32_t)P;
    d.npad = (int32_t)round_up(std::max<int64_t>(n, 1), dla::KC);
    d.min_z = min_z_dla[q];
    d.max_z = max_z_dla[q];
    d.log_prior_dla = log_prior_dla[q];
    d.log_prior_no_dla = log_prior_no_dla[q];
    const int32_t* pi = profile_index + pix_ptr[q];
    bool ident = true;
    for (int64_t j = 0; j < n; ++j) {
      if (pi[j] < 0 || pi[j] > P - 7 || (j > 0 && pi[j] <= pi[j - 1]))
        BAD_ARG(c, "dla_stage_batch: profile_index must be strictly increasing within [0, n_full)");
      ident = ident && (pi[j] == j);
    }
    d.ident = ident ? 1 : 0;
  }
  // chunk plan: W+U of a chunk fit the workspace budget; at least a few chunks when the batch is
  // large enough, so that copies of one chunk overlap the kernels of another (dla_eval_batch)
  c->chunks.clear();
  // exact-slice contraction: eligible batches (three lines, any rank) plan their chunks for either kernel family; a
  // chunk takes the slice path when all of its quasars use the as-written gather
  namespace sl = dla::sl;
  const bool slices_elig = c->use_slices && c->p.k <= dla::KGEN_MAX && c->p.num_lines == 3;
  const int64_t sl_gcols = sl::slice_layout(c->p.k).gcols, sl_bqstep = (int64_t)sl::B_COLSTEP * sl_gcols;
  const int64_t S_pad128 = round_up(S + 1, sl::SB), nsb128 = S_pad128 / sl::SB;
  c->h_sq.assign((size_t)Q, SliceQuasar{});
  auto nks_of = [](const QuasarDesc& d) { return (int64_t)((std::max<int32_t>(d.n, 1) + sl::KS - 1) / sl::KS); };
  auto quasar_bytes = [&](const QuasarDesc& d, long double gq) -> long double {
    long double b = 16.0L * S_pad * d.npad + gq;
    if (slices_elig) {
      // sample planes of W and U + table planes + partial sums, once per set; G once
      const long double per_set = 2.0L * nsb128 * nks_of(d) * sl::A_STEP7 + (long double)nks_of(d) * sl_bqstep +
                                  16.0L * dla::K1S_NSEG_MAX * S_pad128;
      b = std::max(b, (c->overlap ? 2 : 1) * per_set + 8.0L * S_pad128 * sl_gcols);
    }
    return b;
  };
  {
    long double total = 0;
    const long double gq = col_geom(c->p.k).generic ? 8.0L * S_pad * col_geom(c->p.k).ncol : 0.0L;  // G rows
    for (int64_t q = 0; q < Q; ++q) total += quasar_bytes(c->h_desc[(size_t)q], gq);
    const int64_t budget = c->p.workspace_bytes;
    int64_t pipe = std::min<int64_t>(max_pipe_chunks, Q / 8);
    // overlapped slice path: the contraction of chunk i runs under the Voigt kernel of chunk i + 1, so a batch wants
    // many chunks (the first Voigt and the last contraction run alone) of at least a few quasars each
    if (slices_elig && c->overlap) pipe = std::max<int64_t>(pipe, std::min<int64_t>(c->overlap_chunks, Q / 6));
    const int64_t nch = std::max<int64_t>(std::max<int64_t>(1, pipe), (int64_t)ceill(total / budget));
    const long double target = total / nch;
    // quasar-count cap of a chunk: the persistent contraction runs nq * nsb tiles in waves of num_sms; among the
    // counts within 1/8 below what the budget allows, take the one whose last wave is fullest (up to 1 % of K2)
    int64_t nq_cap = 65535;   // the per-quasar kernels take the quasar index as gridDim.y
    if (Q > 0) {
      long double max_bts = 0;
      for (int64_t q = 0; q < Q; ++q) max_bts = std::max(max_bts, quasar_bytes(c->h_desc[(size_t)q], gq));
      const int64_t nq_budget = std::max<int64_t>(1, (int64_t)(budget / max_bts));
      if (nq_budget >= 8 && nq_budget < Q) {
        const int64_t nsb = S_pad / dla::TS;
        double best = -1.0;
        for (int64_t cand = nq_budget; cand >= nq_budget - nq_budget / 8; --cand) {
          const int64_t tiles = cand * nsb, waves = (tiles + c->num_sms - 1) / c->num_sms;
          const double eff = (double)tiles / (double)(waves * c->num_sms);
          if (eff > best + 1e-9) { best = eff; nq_cap = cand; }
        }
      }
    }
    int64_t q = 0;
    while (q < Q) {
      Chunk ch{(int)q, 0, 0, 0, 0, 0, true, 0, 0, 0, false};
      long double acc = 0;
      while (q < Q) {
        QuasarDesc& d = c->h_desc[(size_t)q];
        const long double bts = quasar_bytes(d, gq);
        if (ch.nq > 0 && (acc + bts > budget || acc >= target || ch.nq >= nq_cap)) break;
        d.wu_off = ch.wu_doubles;
        d.bmat_off = ch.bmat_doubles;
        ch.wu_doubles += S_pad * d.npad;
        ch.bmat_doubles += (int64_t)d.npad * dla::NCOL;
        ch.max_npad = std::max(ch.max_npad, (int)d.npad);
        ch.max_P = std::max(ch.max_P, (int)d.P);
        ch.all_ident = ch.all_ident && d.ident != 0;
        SliceQuasar& sq = c->h_sq[(size_t)q];
        sq.nks = (int32_t)nks_of(d);
        sq.a_off = ch.a_bytes;
        sq.b_off = ch.b_bytes;
        ch.a_bytes += nsb128 * sq.nks * sl::A_STEP7;
        ch.b_bytes += (int64_t)sq.nks * sl_bqstep;
        ch.max_nks = std::max(ch.max_nks, (int)sq.nks);
        acc += bts;
        ++ch.nq;
        ++q;
      }
      // int32 accumulators hold 7 * 2^15 * pixels: far beyond any spectrum of the pipeline, checked all the same
      ch.slices = slices_elig && ch.all_ident && ch.max_nks * sl::KS <= 8192;
      c->chunks.push_back(ch);
    }
  }
  const int64_t npix = pix_ptr[Q], npadtot = pad_ptr[Q];
  const int k = c->p.k;
  auto need = [&](DevBuf& b, size_t bytes) { return b.ensure(std::max<size_t>(bytes, 32)); };
  CU_TRY(c, need(c->d_desc, (size_t)Q * sizeof(QuasarDesc)));
  CU_TRY(c, need(c->d_padded, (size_t)npadtot * 8));
  CU_TRY(c, need(c->d_pidx, (size_t)npix * 4));
  CU_TRY(c, need(c->d_krank, (size_t)npadtot * 4));
  CU_TRY(c, need(c->d_pix4, (size_t)npix * 32));
  CU_TRY(c, need(c->d_sq, (size_t)Q * sizeof(SliceQuasar)));
  if (slices_elig) {
    CU_TRY(c, need(c->d_isc, (size_t)npix * 16));
    CU_TRY(c, need(c->d_csc, (size_t)npix * 16));
  }
  CU_TRY(c, need(c->d_flux, (size_t)npix * 8));
  CU_TRY(c, need(c->d_mu, (size_t)npix * 8));
  CU_TRY(c, need(c->d_om, (size_t)npix * 8));
  CU_TRY(c, need(c->d_nv, (size_t)npix * 8));
  CU_TRY(c, need(c->d_M, (size_t)npix * k * 8));
  const size_t q8 = (size_t)std::max<int64_t>(Q, 1) * 8;
  if (c->p.store_sample_ll) CU_TRY(c, need(c->d_sample_ll, (size_t)Q * S * 8));
  CU_TRY(c, c->d_ll_no.ensure(q8));
  CU_TRY(c, c->d_ll_dla.ensure(q8));
  CU_TRY(c, c->d_lp_no.ensure(q8));
  CU_TRY(c, c->d_lp_dla.ensure(q8));
  CU_TRY(c, c->d_post.ensure(2 * q8));
  CU_TRY(c, c->d_map.ensure(q8));
  {
    int64_t wu = 0, bm = 0, ab = 0, bb = 0;
    int nq = 0, nqs = 0;
    for (const Chunk& ch : c->chunks) {
      nq = std::max(nq, ch.nq);
      if (!ch.slices) {
        wu = std::max(wu, ch.wu_doubles);
        bm = std::max(bm, ch.bmat_doubles);
      } else {
        ab = std::max(ab, ch.a_bytes);
        bb = std::max(bb, ch.b_bytes);
        nqs = std::max(nqs, ch.nq);
      }
    }
    int rc = ensure_workspace(c, c->p.k, std::max<int64_t>(wu, 2), std::max<int64_t>(bm, 2), std::max(nq, 1), S,
                              !c->p.store_sample_ll, ab, bb, nqs);
    if (rc != DLA_OK) return rc;
  }
  int rc = ensure_events(c, std::max<size_t>(1, c->chunks.size()));
  if (rc != DLA_OK) return rc;
  c->hin = {pix_ptr, pad_ptr, padded_wavelengths, flux, mu, omega2, noise_variance, M, profile_index, nullptr, nullptr};
  c->Q = Q;
  c->timing = dla_timing{};
  return DLA_OK;
}

// H2D of the inputs of chunk i on stream st (chunk 0 also carries the descriptors of the whole
// batch); records the chunk's "inputs resident" event.
static int upload_chunk(dla_ctx* c, size_t i, cudaStream_t st) {
  NvtxRange r("dla H2D chunk inputs");
  const Chunk& ch = c->chunks[i];
  const auto& h = c->hin;
  const int k = c->p.k;
  auto up = [&](DevBuf& b, const void* src, size_t off_bytes, size_t bytes) -> cudaError_t {
    if (bytes == 0) return cudaSuccess;
    c->timing.h2d_bytes += (int64_t)bytes;
    return cudaMemcpyAsync((char*)b.p + off_bytes, (const char*)src + off_bytes, bytes, cudaMemcpyHostToDevice, st);
  };
  if (i == 0) {
    CU_TRY(c, up(c->d_desc, c->h_desc.data(), 0, (size_t)c->Q * sizeof(QuasarDesc)));
    CU_TRY(c, up(c->d_sq, c->h_sq.data(), 0, (size_t)c->Q * sizeof(SliceQuasar)));
  }
  const int64_t p0 = h.pix_ptr[ch.q0], p1 = h.pix_ptr[ch.q0 + ch.nq];
  const int64_t w0 = h.pad_ptr[ch.q0], w1 = h.pad_ptr[ch.q0 + ch.nq];
  CU_TRY(c, up(c->d_padded, h.padded, (size_t)w0 * 8, (size_t)(w1 - w0) * 8));
  CU_TRY(c, up(c->d_pidx, h.pidx, (size_t)p0 * 4, (size_t)(p1 - p0) * 4));
  CU_TRY(c, up(c->d_flux, h.flux, (size_t)p0 * 8, (size_t)(p1 - p0) * 8));
  CU_TRY(c, up(c->d_nv, h.nv, (size_t)p0 * 8, (size_t)(p1 - p0) * 8));
  if (c->dev_prep) {  // mu, omega2, M are interpolated on the device from the resident model
    if (i == 0) CU_TRY(c, up(c->d_zq, h.zq, 0, (size_t)c->Q * 8));
    CU_TRY(c, up(c->d_lamk, h.lamk, (size_t)p0 * 8, (size_t)(p1 - p0) * 8));
  } else {
    CU_TRY(c, up(c->d_mu, h.mu, (size_t)p0 * 8, (size_t)(p1 - p0) * 8));
    CU_TRY(c, up(c->d_om, h.om, (size_t)p0 * 8, (size_t)(p1 - p0) * 8));
    CU_TRY(c, up(c->d_M, h.M, (size_t)p0 * k * 8, (size_t)(p1 - p0) * k * 8));
  }
  CU_TRY(c, cudaEventRecord(c->ev_sync[2 * i], st));
  return DLA_OK;
}

// kernels of chunk i on the compute stream, after its inputs are resident
static int run_chunk(dla_ctx* c, size_t i) {
  NvtxRange r("dla chunk kernels");
  const Chunk& ch = c->chunks[i];
  if (g_verbose)
    fprintf(stderr, "dla_b200: chunk %zu of %zu: quasars %d..%d of %lld, %s contraction\n", i + 1, c->chunks.size(), ch.q0 + 1,
            ch.q0 + ch.nq, (long long)c->Q, ch.slices ? "int8 slice" : "FP64 DMMA");
  const int64_t S = c->p.num_samples;
  Inputs in{c->d_desc.as<QuasarDesc>(), c->d_padded.as<double>(), c->d_flux.as<double>(), c->d_mu.as<double>(),
            c->d_om.as<double>(), c->d_nv.as<double>(), c->d_M.as<double>(), c->d_pidx.as<int32_t>(),
            c->d_pix4.as<double2>(), c->d_sq.as<SliceQuasar>()};
  CU_TRY(c, cudaStreamWaitEvent(c->stream, c->ev_sync[2 * i], 0));
  if (c->dev_prep) {
    const dla_prep_params& pp = c->prep;
    dla::PrepConsts pc{pp.lya_wavelength, exp(pp.log_c_0), exp(pp.log_tau_0), exp(pp.log_beta)};  // process_qsos.m:84-86
    dim3 grid((unsigned)((ch.max_npad + 127) / 128), (unsigned)ch.nq);
    dla::prep_model_kernel<<<grid, 128, 0, c->stream>>>(
        in.desc, ch.q0, c->d_zq.as<double>(), c->d_lamk.as<double>(), (int)c->n_rest, c->d_mrest.as<double>(),
        c->d_mmu.as<double>(), c->d_mM.as<double>(), c->d_mlo.as<double>(), pc, c->p.k, c->d_mu.as<double>(),
        c->d_om.as<double>(), c->d_M.as<double>());
    c->timing.launches += 1;
  }
  double* ll_chunk = c->p.store_sample_ll ? c->d_sample_ll.as<double>() + (int64_t)ch.q0 * S : c->d_ll_chunk.as<double>();
  int rc = launch_chunk(c, in, ch, S, c->p.k, c->p.num_lines, c->d_offset.as<double>(), c->d_nhi.as<double>(),
                        c->d_perm.as<int32_t>(), ll_chunk, c->d_ll_no.as<double>(), c->d_ll_dla.as<double>(),
                        c->d_lp_no.as<double>(), c->d_lp_dla.as<double>(), c->d_post.as<double>(),
                        c->d_map.as<int32_t>(), &c->ev_time[NTEV * i]);
  if (rc != DLA_OK) return rc;
  CU_TRY(c, cudaEventRecord(c->ev_sync[2 * i + 1], (ch.slices && c->overlap) ? c->stream2 : c->stream));
  c->timing.chunks++;
  return DLA_OK;
}

static int dla_stage_batch_impl(dla_ctx* c, int64_t Q, const int64_t* pix_ptr, const int64_t* pad_ptr,
                    const double* padded_wavelengths, const int32_t* profile_index, const double* flux,
                    const double* mu, const double* omega2, const double* noise_variance, const double* M,
                    const double* min_z_dla, const double* max_z_dla, const double* log_prior_dla,
                    const double* log_prior_no_dla) {
  if (!c) return DLA_ERR_BAD_ARG;
  c->dev_prep = false;
  // everything is resident before the first kernel: no reason to split beyond the workspace budget
  int rc = plan_batch(c, Q, pix_ptr, pad_ptr, padded_wavelengths, profile_index, flux, mu, omega2, noise_variance, M,
                      min_z_dla, max_z_dla, log_prior_dla, log_prior_no_dla, 1);
  if (rc != DLA_OK) return rc;
  CU_TRY(c, cudaEventRecord(c->ev[5], c->s_in));
  for (size_t i = 0; i < c->chunks.size(); ++i)
    if ((rc = upload_chunk(c, i, c->s_in)) != DLA_OK) return rc;
  CU_TRY(c, cudaEventRecord(c->ev[6], c->s_in));
  CU_TRY(c, cudaStreamSynchronize(c->s_in));
  c->hin = {};
  float ms;
  CU_TRY(c, cudaEventElapsedTime(&ms, c->ev[5], c->ev[6]));
  c->timing.ms_h2d = ms;
  c->staged = true;
  return DLA_OK;
}

static int dla_run_staged_impl(dla_ctx* c) {
  if (!c) return DLA_ERR_BAD_ARG;
  if (!c->staged) {
    c->err = "dla_run_staged: no batch staged";
    return DLA_ERR_STATE;
  }
  CU_TRY(c, cudaSetDevice(c->device));
  const int64_t keep_h2d = c->timing.h2d_bytes;
  const double keep_ms_h2d = c->timing.ms_h2d;
  c->timing = dla_timing{};
  c->timing.h2d_bytes = keep_h2d;
  c->timing.ms_h2d = keep_ms_h2d;
  for (size_t i = 0; i < c->chunks.size(); ++i) {
    int rc = run_chunk(c, i);
    if (rc != DLA_OK) return rc;
  }
  int rc = sync_compute(c);
  if (rc != DLA_OK) return rc;
  if (!c->trace_path.empty()) {   // the trace of the LAST run (the counter restarts with every run)
    unsigned int n = 0, zero = 0;
    CU_TRY(c, cudaMemcpyFromSymbol(&n, dla::sl::g_trace_n, sizeof n));
    CU_TRY(c, cudaMemcpyToSymbol(dla::sl::g_trace_n, &zero, sizeof zero));
    n = std::min(n, dla::sl::TRACE_CAP);
    std::vector<unsigned long long> h((size_t)n * 4);
    if (n) CU_TRY(c, cudaMemcpy(h.data(), c->d_trace.p, (size_t)n * 32, cudaMemcpyDeviceToHost));
    if (FILE* f = fopen(c->trace_path.c_str(), "wb")) {
      fwrite(h.data(), 8, h.size(), f);
      fclose(f);
    }
  }
  rc = collect_timing(c, c->chunks.size());
  if (rc != DLA_OK) return rc;
  c->ran = true;
  return DLA_OK;
}

// D2H of the per-quasar outputs (everything except sample_ll) on stream st
static int download_small(dla_ctx* c, cudaStream_t st, double* ll_no_dla, double* ll_dla, double* log_post_no_dla,
                          double* log_post_dla, double* model_posteriors, int32_t* map_index) {
  const int64_t Q = c->Q;
  auto down = [&](void* dst, const DevBuf& b, size_t bytes) -> cudaError_t {
    if (!dst || bytes == 0) return cudaSuccess;
    c->timing.d2h_bytes += (int64_t)bytes;
    return cudaMemcpyAsync(dst, b.p, bytes, cudaMemcpyDeviceToHost, st);
  };
  CU_TRY(c, down(ll_no_dla, c->d_ll_no, (size_t)Q * 8));
  CU_TRY(c, down(ll_dla, c->d_ll_dla, (size_t)Q * 8));
  CU_TRY(c, down(log_post_no_dla, c->d_lp_no, (size_t)Q * 8));
  CU_TRY(c, down(log_post_dla, c->d_lp_dla, (size_t)Q * 8));
  CU_TRY(c, down(model_posteriors, c->d_post, (size_t)Q * 16));
  CU_TRY(c, down(map_index, c->d_map, (size_t)Q * 4));
  return DLA_OK;
}

static int dla_fetch_staged_impl(dla_ctx* c, double* sample_ll, double* ll_no_dla, double* ll_dla, double* log_post_no_dla,
                     double* log_post_dla, double* model_posteriors, int32_t* map_index) {
  if (!c) return DLA_ERR_BAD_ARG;
  if (!c->ran) {
    c->err = "dla_fetch_staged: dla_run_staged has not completed";
    return DLA_ERR_STATE;
  }
  if (sample_ll && !c->p.store_sample_ll) BAD_ARG(c, "dla_fetch_staged: sample_ll requested but store_sample_ll = 0");
  CU_TRY(c, cudaSetDevice(c->device));
  const int64_t Q = c->Q, S = c->p.num_samples;
  c->timing.d2h_bytes = 0;
  CU_TRY(c, cudaEventRecord(c->ev[5], c->stream));
  if (sample_ll && Q * S > 0) {
    c->timing.d2h_bytes += Q * S * 8;
    CU_TRY(c, cudaMemcpyAsync(sample_ll, c->d_sample_ll.p, (size_t)Q * S * 8, cudaMemcpyDeviceToHost, c->stream));
  }
  int rc = download_small(c, c->stream, ll_no_dla, ll_dla, log_post_no_dla, log_post_dla, model_posteriors, map_index);
  if (rc != DLA_OK) return rc;
  CU_TRY(c, cudaEventRecord(c->ev[6], c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  float ms;
  CU_TRY(c, cudaEventElapsedTime(&ms, c->ev[5], c->ev[6]));
  c->timing.ms_d2h = ms;
  return DLA_OK;
}

// One call, pipelined over the chunks: H2D of chunk i+1 (stream s_in) and D2H of chunk i-1's
// sample log-likelihoods (stream s_out) run under the kernels of chunk i (compute stream).
static int run_pipelined(dla_ctx* c, double* sample_ll, double* ll_no_dla, double* ll_dla, double* log_post_no_dla,
                         double* log_post_dla, double* model_posteriors, int32_t* map_index) {
  int rc = DLA_OK;
  const int64_t S = c->p.num_samples;
  const size_t nch = c->chunks.size();
  CU_TRY(c, cudaEventRecord(c->ev[5], c->s_in));
  for (size_t i = 0; i < nch; ++i)
    if ((rc = upload_chunk(c, i, c->s_in)) != DLA_OK) return rc;
  CU_TRY(c, cudaEventRecord(c->ev[6], c->s_in));
  for (size_t i = 0; i < nch; ++i) {
    if ((rc = run_chunk(c, i)) != DLA_OK) return rc;
    const Chunk& ch = c->chunks[i];
    if (sample_ll && (int64_t)ch.nq * S > 0) {
      CU_TRY(c, cudaStreamWaitEvent(c->s_out, c->ev_sync[2 * i + 1], 0));
      const size_t off = (size_t)ch.q0 * S * 8, bytes = (size_t)ch.nq * S * 8;
      c->timing.d2h_bytes += (int64_t)bytes;
      CU_TRY(c, cudaMemcpyAsync((char*)sample_ll + off, (const char*)c->d_sample_ll.p + off, bytes,
                                cudaMemcpyDeviceToHost, c->s_out));
    }
  }
  if (nch > 0) CU_TRY(c, cudaStreamWaitEvent(c->s_out, c->ev_sync[2 * (nch - 1) + 1], 0));
  rc = download_small(c, c->s_out, ll_no_dla, ll_dla, log_post_no_dla, log_post_dla, model_posteriors, map_index);
  if (rc != DLA_OK) return rc;
  CU_TRY(c, cudaStreamSynchronize(c->s_in));
  if ((rc = sync_compute(c)) != DLA_OK) return rc;
  CU_TRY(c, cudaStreamSynchronize(c->s_out));
  c->hin = {};
  float ms;
  CU_TRY(c, cudaEventElapsedTime(&ms, c->ev[5], c->ev[6]));
  c->timing.ms_h2d = ms;  // span of the input copies on their own stream (they overlap the kernels)
  if ((rc = collect_timing(c, nch)) != DLA_OK) return rc;
  c->staged = true;
  c->ran = true;
  return DLA_OK;
}

static int dla_eval_batch_impl(dla_ctx* c, int64_t Q, const int64_t* pix_ptr, const int64_t* pad_ptr,
                   const double* padded_wavelengths, const int32_t* profile_index, const double* flux,
                   const double* mu, const double* omega2, const double* noise_variance, const double* M,
                   const double* min_z_dla, const double* max_z_dla, const double* log_prior_dla,
                   const double* log_prior_no_dla, double* sample_ll, double* ll_no_dla, double* ll_dla,
                   double* log_post_no_dla, double* log_post_dla, double* model_posteriors, int32_t* map_index) {
  if (!c) return DLA_ERR_BAD_ARG;
  if (sample_ll && !c->p.store_sample_ll) BAD_ARG(c, "dla_eval_batch: sample_ll requested but store_sample_ll = 0");
  c->dev_prep = false;
  int rc = plan_batch(c, Q, pix_ptr, pad_ptr, padded_wavelengths, profile_index, flux, mu, omega2, noise_variance, M,
                      min_z_dla, max_z_dla, log_prior_dla, log_prior_no_dla, 4);
  if (rc != DLA_OK) return rc;
  return run_pipelined(c, sample_ll, ll_no_dla, ll_dla, log_post_no_dla, log_post_dla, model_posteriors, map_index);
}

static int dla_set_model_impl(dla_ctx* c, int64_t n_rest, const double* rest_wavelengths, const double* mu, const double* M,
                  const double* log_omega, const dla_prep_params* prep) {
  if (!c) return DLA_ERR_BAD_ARG;
  if (!rest_wavelengths || !mu || !M || !log_omega || !prep) BAD_ARG(c, "dla_set_model: NULL pointer");
  if (n_rest < 2 || n_rest > (1 << 24)) BAD_ARG(c, "dla_set_model: n_rest out of range");
  if (prep->width != 3) BAD_ARG(c, "dla_set_model: width must be 3 (the instrument profile of voigt.c:229 has 7 taps)");
  for (int64_t i = 1; i < n_rest; ++i)
    if (!(rest_wavelengths[i] > rest_wavelengths[i - 1])) BAD_ARG(c, "dla_set_model: rest_wavelengths must be strictly increasing");
  CU_TRY(c, cudaSetDevice(c->device));
  const int k = c->p.k;
  {
    // the same learned model with every call (the MEX gateways): nothing to validate or copy again
    const size_t nr = (size_t)n_rest, nb = nr * 8;
    const std::vector<double>& h = c->h_set_model;
    if (c->have_model && c->n_rest == n_rest && h.size() == nr * (size_t)(k + 3) &&
        memcmp(&c->prep, prep, sizeof *prep) == 0 && memcmp(h.data(), rest_wavelengths, nb) == 0 &&
        memcmp(h.data() + nr, mu, nb) == 0 && memcmp(h.data() + 2 * nr, log_omega, nb) == 0 &&
        memcmp(h.data() + 3 * nr, M, nb * (size_t)k) == 0)
      return DLA_OK;
    c->have_model = false;
  }
  CU_TRY(c, c->d_mrest.ensure((size_t)n_rest * 8));
  CU_TRY(c, c->d_mmu.ensure((size_t)n_rest * 8));
  CU_TRY(c, c->d_mlo.ensure((size_t)n_rest * 8));
  CU_TRY(c, c->d_mM.ensure((size_t)n_rest * k * 8));
  CU_TRY(c, cudaMemcpyAsync(c->d_mrest.p, rest_wavelengths, (size_t)n_rest * 8, cudaMemcpyHostToDevice, c->stream));
  CU_TRY(c, cudaMemcpyAsync(c->d_mmu.p, mu, (size_t)n_rest * 8, cudaMemcpyHostToDevice, c->stream));
  CU_TRY(c, cudaMemcpyAsync(c->d_mlo.p, log_omega, (size_t)n_rest * 8, cudaMemcpyHostToDevice, c->stream));
  CU_TRY(c, cudaMemcpyAsync(c->d_mM.p, M, (size_t)n_rest * k * 8, cudaMemcpyHostToDevice, c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  c->n_rest = n_rest;
  memset(&c->prep, 0, sizeof c->prep);  // padding bytes too: the unchanged-model test compares the whole struct
  c->prep = *prep;
  {
    std::vector<double>& h = c->h_set_model;
    h.resize((size_t)n_rest * (size_t)(k + 3));
    const size_t nr = (size_t)n_rest;
    memcpy(h.data(), rest_wavelengths, nr * 8);
    memcpy(h.data() + nr, mu, nr * 8);
    memcpy(h.data() + 2 * nr, log_omega, nr * 8);
    memcpy(h.data() + 3 * nr, M, nr * 8 * (size_t)k);
  }
  c->have_model = true;
  return DLA_OK;
}

// One host pass over quasars [qa, qb): process_qsos.m:102-116 (selection), :160-161 (z range), :169-180 (padded
// grid, gather).  Pure host code (no CUDA calls): runs on a helper thread for the NEXT sub-batch.
static void select_spectra(const dla_prep_params pp, dla_ctx::SpectraStage* stp, int64_t qa, int64_t qb,
                           const int64_t* spec_ptr, const double* wavelengths, const double* flux,
                           const double* noise_variance, const uint8_t* pixel_mask, const double* z_qso,
                           int32_t mask_gather) {
  dla_ctx::SpectraStage& st = *stp;
  const int64_t Qs = qb - qa;
  st.rc = DLA_OK;
  st.err = nullptr;
  st.pix_ptr.assign((size_t)Qs + 1, 0);
  st.pad_ptr.assign((size_t)Qs + 1, 0);
  st.minz.assign((size_t)Qs, 0.0);
  st.maxz.assign((size_t)Qs, 0.0);
  double *lamk = st.lamk.as<double>(), *fk = st.fk.as<double>(), *nk = st.nk.as<double>();
  double* padded = st.padded.as<double>();
  int32_t* pidx = st.pidx.as<int32_t>();
  int64_t nkept = 0, npadded = 0;
  for (int64_t q = qa; q < qb; ++q) {
    const int64_t a = spec_ptr[q], b = spec_ptr[q + 1];
    const double opz = 1.0 + z_qso[q];
    const int64_t pad0 = npadded;
    npadded += 3;  // left pads, filled below
    int64_t n_full = 0, n = 0;
    double umin = INFINITY, umax = -INFINITY, kmin = INFINITY, kmax = -INFINITY;
    for (int64_t i = a; i < b; ++i) {
      const double w = wavelengths[i];
      const double rest = w / opz;                                            // :102, set_parameters.m:14-15
      if (!(rest >= pp.min_lambda && rest <= pp.max_lambda)) continue;        // :104-105
      padded[npadded++] = w;                                                  // :109 this_unmasked_wavelengths
      umin = std::min(umin, w); umax = std::max(umax, w);
      if (!pixel_mask[i]) {                                                   // :111-116
        lamk[nkept] = w; fk[nkept] = flux[i]; nk[nkept] = noise_variance[i];
        pidx[nkept] = (int32_t)(mask_gather ? n_full : n);                    // :180/:189 as written, or the unmasked positions
        ++nkept;
        kmin = std::min(kmin, w); kmax = std::max(kmax, w);
        ++n;
      }
      ++n_full;
    }
    if (n_full == 0 || n == 0) {
      st.rc = DLA_ERR_BAD_ARG;
      st.err = "dla_eval_spectra: a quasar has no unmasked pixel in the modelled range";
      return;
    }
    // :169-177: width log-spaced pads each side, 10.^linspace(...)
    const double lo = log10(umin), hi = log10(umax), ps = pp.pixel_spacing;
    auto linspace3 = [](double x0, double x1, double* out) {  // numpy/MATLAB linspace(x0, x1, 3)
      const double step = (x1 - x0) / 2.0;
      out[0] = x0; out[1] = x0 + step; out[2] = x1;
    };
    double e[3];
    linspace3(lo - 3 * ps, lo - ps, e);
    for (int t = 0; t < 3; ++t) padded[pad0 + t] = pow(10.0, e[t]);
    linspace3(hi + ps, hi + 3 * ps, e);
    for (int t = 0; t < 3; ++t) padded[npadded++] = pow(10.0, e[t]);
    // :160-161, set_parameters.m:66-73 (on the kept wavelengths)
    const size_t ql = (size_t)(q - qa);
    st.maxz[ql] = (kmax / pp.lya_wavelength - 1) - pp.max_z_cut;
    st.minz[ql] = std::max(kmin / pp.lya_wavelength - 1,
                           pp.lyman_limit * (1 + z_qso[q]) / pp.lya_wavelength - 1 + pp.min_z_cut);
    st.pix_ptr[ql + 1] = st.pix_ptr[ql] + n;
    st.pad_ptr[ql + 1] = st.pad_ptr[ql] + n_full + 6;
  }
  st.nkept = nkept;
}

static int dla_eval_spectra_impl(dla_ctx* c, int64_t Q, const int64_t* spec_ptr, const double* wavelengths, const double* flux,
                     const double* noise_variance, const uint8_t* pixel_mask, const double* z_qso,
                     const double* log_prior_dla, const double* log_prior_no_dla, int32_t mask_gather,
                     double* sample_ll, double* ll_no_dla, double* ll_dla, double* log_post_no_dla,
                     double* log_post_dla, double* model_posteriors, int32_t* map_index, double* min_z_out,
                     double* max_z_out) {
  if (!c) return DLA_ERR_BAD_ARG;
  if (!c->have_model) {
    c->err = "dla_eval_spectra: dla_set_model has not been called";
    return DLA_ERR_STATE;
  }
  if (Q < 0 || !spec_ptr || !wavelengths || !flux || !noise_variance || !pixel_mask || !z_qso || !log_prior_dla ||
      !log_prior_no_dla)
    BAD_ARG(c, "dla_eval_spectra: NULL pointer");
  if (mask_gather != 0 && mask_gather != 1) BAD_ARG(c, "dla_eval_spectra: mask_gather must be 0 or 1");
  if (sample_ll && !c->p.store_sample_ll) BAD_ARG(c, "dla_eval_spectra: sample_ll requested but store_sample_ll = 0");
  for (int64_t q = 0; q < Q; ++q)
    if (spec_ptr[q + 1] < spec_ptr[q]) BAD_ARG(c, "dla_eval_spectra: spec_ptr must be non-decreasing");
  CU_TRY(c, cudaSetDevice(c->device));
  // Sub-batches of <= SUBQ quasars: the selection pass of sub-batch i+1 runs on a helper thread while the device
  // evaluates sub-batch i, and the staging / device input buffers stay bounded however large Q is.
  const int64_t SUBQ = 2048;
  const int64_t nsub = std::max<int64_t>(1, (Q + SUBQ - 1) / SUBQ);
  int64_t maxtot = 0, maxq = 0;
  for (int64_t i = 0; i < nsub; ++i) {
    const int64_t qa = i * SUBQ, qb = std::min(Q, qa + SUBQ);
    maxtot = std::max(maxtot, spec_ptr[qb] - spec_ptr[qa]);
    maxq = std::max(maxq, qb - qa);
  }
  for (int b = 0; b < (nsub > 1 ? 2 : 1); ++b) {
    auto& st = c->sstage[b];
    CU_TRY(c, st.lamk.ensure((size_t)maxtot * 8 + 64));
    CU_TRY(c, st.fk.ensure((size_t)maxtot * 8 + 64));
    CU_TRY(c, st.nk.ensure((size_t)maxtot * 8 + 64));
    CU_TRY(c, st.pidx.ensure((size_t)maxtot * 4 + 64));
    CU_TRY(c, st.padded.ensure(((size_t)maxtot + 6 * (size_t)maxq) * 8 + 64));
  }
  const dla_prep_params pp = c->prep;
  select_spectra(pp, &c->sstage[0], 0, std::min(Q, SUBQ), spec_ptr, wavelengths, flux, noise_variance, pixel_mask,
                 z_qso, mask_gather);
  dla_timing acc{};
  int rc = DLA_OK;
  for (int64_t i = 0; i < nsub && rc == DLA_OK; ++i) {
    const int64_t qa = i * SUBQ, qb = std::min(Q, qa + SUBQ), Qs = qb - qa;
    auto& st = c->sstage[i & 1];
    std::thread next;
    if (i + 1 < nsub)
      next = std::thread(select_spectra, pp, &c->sstage[(i + 1) & 1], qb, std::min(Q, qb + SUBQ), spec_ptr, wavelengths,
                         flux, noise_variance, pixel_mask, z_qso, mask_gather);
    if (st.rc != DLA_OK) {
      c->err = st.err ? st.err : "dla_eval_spectra: selection failed";
      rc = st.rc;
    }
    if (rc == DLA_OK) {
      c->dev_prep = true;
      // mu / omega2 / M are produced on the device: any non-NULL pointer satisfies the planner
      const double* fk = st.fk.as<double>();
      const double* dummy = fk;
      rc = plan_batch(c, Qs, st.pix_ptr.data(), st.pad_ptr.data(), st.padded.as<double>(), st.pidx.as<int32_t>(), fk,
                      dummy, dummy, st.nk.as<double>(), dummy, st.minz.data(), st.maxz.data(), log_prior_dla + qa,
                      log_prior_no_dla + qa, 4);
      if (rc == DLA_OK) {
        c->hin.lamk = st.lamk.as<double>();
        c->hin.zq = z_qso + qa;
        cudaError_t e1 = c->d_lamk.ensure(std::max<size_t>((size_t)st.nkept * 8, 32));
        cudaError_t e2 = c->d_zq.ensure(std::max<size_t>((size_t)Qs * 8, 32));
        if (e1 != cudaSuccess || e2 != cudaSuccess) {
          c->err = "dla_eval_spectra: device allocation failed";
          rc = DLA_ERR_OOM;
        }
      }
      const int64_t S = c->p.num_samples;
      if (rc == DLA_OK)
        rc = run_pipelined(c, sample_ll ? sample_ll + qa * S : nullptr, ll_no_dla ? ll_no_dla + qa : nullptr,
                           ll_dla ? ll_dla + qa : nullptr, log_post_no_dla ? log_post_no_dla + qa : nullptr,
                           log_post_dla ? log_post_dla + qa : nullptr,
                           model_posteriors ? model_posteriors + 2 * qa : nullptr, map_index ? map_index + qa : nullptr);
      c->dev_prep = false;
      c->staged = false;  // the staged-run API needs host-prepared inputs
      if (rc == DLA_OK) {
        if (min_z_out) memcpy(min_z_out + qa, st.minz.data(), (size_t)Qs * 8);
        if (max_z_out) memcpy(max_z_out + qa, st.maxz.data(), (size_t)Qs * 8);
        const dla_timing& t = c->timing;
        acc.ms_phi += t.ms_phi; acc.ms_voigt += t.ms_voigt; acc.ms_contract += t.ms_contract;
        acc.ms_reduce += t.ms_reduce; acc.ms_total += t.ms_total; acc.ms_h2d += t.ms_h2d; acc.ms_d2h += t.ms_d2h;
        acc.ms_chol += t.ms_chol; acc.contraction = std::max(acc.contraction, t.contraction);
        acc.h2d_bytes += t.h2d_bytes; acc.d2h_bytes += t.d2h_bytes; acc.chunks += t.chunks; acc.launches += t.launches;
      }
    }
    if (next.joinable()) next.join();
  }
  if (rc != DLA_OK) return rc;
  c->timing = acc;
  return DLA_OK;
}

static int dla_train_set_data_impl(dla_ctx* c, int64_t Q, int64_t n, const double* y, const double* l1pz, const double* nv) {
  if (!c) return DLA_ERR_BAD_ARG;
  if (!y || !l1pz || !nv) BAD_ARG(c, "dla_train_set_data: NULL pointer");
  if (Q < 1 || n < 1 || n > (1 << 20) || Q > (1 << 26)) BAD_ARG(c, "dla_train_set_data: sizes out of range");
  if (c->p.k > dla::TRAIN_KMAX) BAD_ARG(c, "dla_train_set_data: the training objective supports k <= 32");
  CU_TRY(c, cudaSetDevice(c->device));
  const size_t bytes = (size_t)Q * n * 8;
  CU_TRY(c, c->t_y.ensure(bytes));
  CU_TRY(c, c->t_l.ensure(bytes));
  CU_TRY(c, c->t_nv.ensure(bytes));
  CU_TRY(c, cudaMemcpyAsync(c->t_y.p, y, bytes, cudaMemcpyHostToDevice, c->stream));
  CU_TRY(c, cudaMemcpyAsync(c->t_l.p, l1pz, bytes, cudaMemcpyHostToDevice, c->stream));
  CU_TRY(c, cudaMemcpyAsync(c->t_nv.p, nv, bytes, cudaMemcpyHostToDevice, c->stream));
  CU_TRY(c, cudaStreamSynchronize(c->stream));
  c->tQ = Q;
  c->tN = n;
  c->staged = false;  // shares the W/U/G workspace with the evaluation path
  c->ran = false;
  return DLA_OK;
}

static int dla_train_objective_impl(dla_ctx* c, const double* x, int32_t apply_priors, double* f, double* g) {
  if (!c) return DLA_ERR_BAD_ARG;
  if (!x || !f || !g) BAD_ARG(c, "dla_train_objective: NULL pointer");
  if (c->tQ == 0) {
    c->err = "dla_train_objective: dla_train_set_data has not been called";
    return DLA_ERR_STATE;
  }
  CU_TRY(c, cudaSetDevice(c->device));
  c->staged = false;
  c->ran = false;
  const int64_t Q = c->tQ;
  const int n = (int)c->tN, k = c->p.k;
  const int T = k * (k + 1) / 2;
  const int npad = (int)round_up(n, dla::KC);
  const int nqb = (int)((Q + dla::TS - 1) / dla::TS);
  const int64_t Qpad = (int64_t)nqb * dla::TS;
  const ColGeom gen = col_geom(k);
  const int64_t nx = (int64_t)n * (k + 1) + 3;
  const int nsplit = (int)std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(32, Q), (600 + (n + 7) / 8 - 1) / ((n + 7) / 8)));
  cudaStream_t st = c->stream;
  CU_TRY(c, c->t_x.ensure((size_t)nx * 8));
  CU_TRY(c, c->t_om2.ensure((size_t)n * 8));
  CU_TRY(c, c->d_W.ensure((size_t)Qpad * npad * 8));
  CU_TRY(c, c->d_U.ensure((size_t)Qpad * npad * 8));
  CU_TRY(c, c->d_bmat.ensure((size_t)npad * gen.ncol * 8));
  CU_TRY(c, c->d_G.ensure((size_t)Qpad * gen.ncol * 8));
  CU_TRY(c, c->t_aux.ensure((size_t)3 * dla::TRAIN_NSEG * Qpad * 8));
  CU_TRY(c, c->t_E.ensure((size_t)dla::TRAIN_NE * Q * n * 8));
  CU_TRY(c, c->t_fq.ensure((size_t)Q * 8));
  CU_TRY(c, c->t_beta.ensure((size_t)Q * k * 8));
  CU_TRY(c, c->t_P.ensure((size_t)Q * T * 8));
  CU_TRY(c, c->t_A1.ensure((size_t)nsplit * n * k * 8));
  CU_TRY(c, c->t_S.ensure((size_t)nsplit * n * T * 8));
  CU_TRY(c, c->t_gp.ensure((size_t)nsplit * n * 4 * 8));
  CU_TRY(c, c->t_g.ensure((size_t)n * (k + 1) * 8));
  CU_TRY(c, c->t_gsc.ensure((size_t)n * 3 * 8));
  CU_TRY(c, c->t_desc.ensure(sizeof(QuasarDesc)));
  QuasarDesc hd{};
  hd.n = n;
  hd.P = n + 6;
  hd.npad = npad;
  hd.ident = 1;
  CU_TRY(c, cudaMemcpyAsync(c->t_desc.p, &hd, sizeof hd, cudaMemcpyHostToDevice, st));
  CU_TRY(c, cudaMemcpyAsync(c->t_x.p, x, (size_t)nx * 8, cudaMemcpyHostToDevice, st));
  const double* dx = c->t_x.as<double>();
  const dla::TrainConsts tc{exp(x[nx - 3]), exp(x[nx - 2]), exp(x[nx - 1])};   // objective.m:30-32
  dla::train_omega2_kernel<<<(n + 255) / 256, 256, 0, st>>>(dx + (int64_t)n * k, n, c->t_om2.as<double>());
  dla::train_weights_kernel<<<nqb, dla::TS * dla::TRAIN_NSEG, 0, st>>>(
      Q, n, npad, c->t_y.as<double>(), c->t_l.as<double>(), c->t_nv.as<double>(), c->t_om2.as<double>(), tc,
      c->d_W.as<double>(), c->d_U.as<double>(), c->t_aux.as<double>(), c->t_E.as<double>());
  {
    dim3 grid((unsigned)(((int64_t)(npad / 4) * gen.ncol + 255) / 256), 1);
    dla::build_bmat_kernel<<<grid, 256, 0, st>>>(c->t_desc.as<QuasarDesc>(), 0, dx, c->d_bmat.as<double>(), k, gen.ncol, gen.voff);
  }
  {
    const size_t smem = (size_t)dla::K2_STAGES * dla::STAGE_DOUBLES * 8 + 64;
    dim3 grid((unsigned)nqb, (unsigned)gen.nblocks);
    launch_contract_block(gen.tpw, grid, smem, st, (const QuasarDesc*)c->t_desc.as<QuasarDesc>(), 0, nqb,
                          (const double*)c->d_W.as<double>(), (const double*)c->d_U.as<double>(),
                          (const double*)c->d_bmat.as<double>(), c->d_G.as<double>(), gen.ncol, gen.voff / 8);
  }
  {
    const size_t smem = (size_t)dla::TRC_WARPS * ((k + 1) * (k + 2) + k * (k + 1) + k) * 8;
    dla::train_chol_kernel<<<(unsigned)((Q + dla::TRC_WARPS - 1) / dla::TRC_WARPS), dla::TRC_WARPS * 32, smem, st>>>(
        Q, Qpad, c->d_G.as<double>(), c->t_aux.as<double>(), k, gen.ncol, gen.voff, c->t_fq.as<double>(),
        c->t_beta.as<double>(), c->t_P.as<double>());
  }
  {
    dim3 grid((unsigned)((n + dla::TRG_WARPS - 1) / dla::TRG_WARPS), (unsigned)nsplit);
    dla::train_grad_kernel<<<grid, dla::TRG_WARPS * 32, 0, st>>>(
        Q, n, k, dx, c->t_E.as<double>(), c->t_beta.as<double>(), c->t_P.as<double>(), tc, nsplit,
        c->t_A1.as<double>(), c->t_S.as<double>(), c->t_gp.as<double>());
  }
  dla::train_finish_kernel<<<(n * (k + 1) + 255) / 256, 256, 0, st>>>(
      n, k, nsplit, dx, c->t_A1.as<double>(), c->t_S.as<double>(), c->t_gp.as<double>(), c->t_g.as<double>(),
      c->t_gsc.as<double>());
  CU_TRY(c, cudaGetLastError());
  std::vector<double> fq((size_t)Q), gsc((size_t)n * 3);
  CU_TRY(c, cudaMemcpyAsync(g, c->t_g.p, (size_t)n * (k + 1) * 8, cudaMemcpyDeviceToHost, st));
  CU_TRY(c, cudaMemcpyAsync(fq.data(), c->t_fq.p, (size_t)Q * 8, cudaMemcpyDeviceToHost, st));
  CU_TRY(c, cudaMemcpyAsync(gsc.data(), c->t_gsc.p, (size_t)n * 3 * 8, cudaMemcpyDeviceToHost, st));
  CU_TRY(c, cudaStreamSynchronize(st));
  double fs = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
  for (int64_t q = 0; q < Q; ++q) fs += fq[(size_t)q];                       // objective.m:50
  for (int p = 0; p < n; ++p) { s1 += gsc[(size_t)p * 3]; s2 += gsc[(size_t)p * 3 + 1]; s3 += gsc[(size_t)p * 3 + 2]; }
  if (apply_priors) {
    s2 += tc.tau_0 * (tc.tau_0 - 0.0023) / (0.0007 * 0.0007);                 // objective.m:60-64
    s3 += tc.beta * (tc.beta - 3.65) / (0.21 * 0.21);                         // :67-71
  }
  *f = fs;
  g[nx - 3] = s1;
  g[nx - 2] = s2;
  g[nx - 1] = s3;
  c->timing.launches = 7;
  return DLA_OK;
}

int dla_get_timing(const dla_ctx* c, dla_timing* out) {
  if (!c || !out) return DLA_ERR_BAD_ARG;
  *out = c->timing;
  return DLA_OK;
}

static int dla_voigt_impl(dla_ctx* c, const double* lambdas, int64_t num_points, double z, double N, int32_t num_lines,
              double* profile) {
  if (!c) return DLA_ERR_BAD_ARG;
  if (!lambdas || !profile) BAD_ARG(c, "dla_voigt: NULL pointer");
  if (num_points < 7 || num_points > (1 << 26)) BAD_ARG(c, "dla_voigt: need at least 7 points (voigt.c:271)");
  if (num_lines <= 0 || num_lines > DLA_MAX_LINES) num_lines = DLA_MAX_LINES;  // voigt.c:266
  CU_TRY(c, cudaSetDevice(c->device));
  DevBuf &lam = c->seam[8], &raw = c->seam[9], &prof = c->seam[10];
  int rc = DLA_OK;
  auto fail = [&](cudaError_t e, const char* what) {
    c->err = std::string("dla_voigt: ") + what + ": " + cudaGetErrorString(e);
    rc = e == cudaErrorMemoryAllocation ? DLA_ERR_OOM : DLA_ERR_CUDA;
  };
  cudaError_t e;
  if ((e = lam.ensure((size_t)num_points * 8)) != cudaSuccess) fail(e, "alloc");
  if (rc == DLA_OK && (e = raw.ensure((size_t)num_points * 8)) != cudaSuccess) fail(e, "alloc");
  if (rc == DLA_OK && (e = prof.ensure((size_t)num_points * 8)) != cudaSuccess) fail(e, "alloc");
  if (rc == DLA_OK && (e = cudaMemcpyAsync(lam.p, lambdas, (size_t)num_points * 8, cudaMemcpyHostToDevice, c->stream)) != cudaSuccess)
    fail(e, "h2d");
  if (rc == DLA_OK) {
    dla::voigt_single_kernel<<<1, 1024, 0, c->stream>>>(lam.as<double>(), (int)num_points, z, N, num_lines,
                                                          raw.as<double>(), prof.as<double>());
    if ((e = cudaGetLastError()) != cudaSuccess) fail(e, "launch");
  }
  if (rc == DLA_OK && (e = cudaMemcpyAsync(profile, prof.p, (size_t)(num_points - 6) * 8, cudaMemcpyDeviceToHost, c->stream)) != cudaSuccess)
    fail(e, "d2h");
  if (rc == DLA_OK && (e = cudaStreamSynchronize(c->stream)) != cudaSuccess) fail(e, "sync");
  if (rc != DLA_OK) cudaStreamSynchronize(c->stream);  // nothing may still read the caller's array
  return rc;
}

// dla_voigt without any allocation: the same kernel on the file-scope scratch.  cudaMemcpyDefault: `lambdas` and
// `profile` may be host or device pointers.
static int dla_voigt_noalloc_impl(dla_ctx* c, const double* lambdas, int64_t num_points, double z, double N, int32_t num_lines,
                                  double* profile) {
  if (!c) return DLA_ERR_BAD_ARG;
  if (!lambdas || !profile) BAD_ARG(c, "dla_voigt_noalloc: NULL pointer");
  if (num_points < 7) BAD_ARG(c, "dla_voigt_noalloc: need at least 7 points (voigt.c:271)");
  if (num_points > DLA_NOALLOC_MAX_POINTS)
    BAD_ARG(c, "dla_voigt_noalloc: num_points exceeds DLA_NOALLOC_MAX_POINTS (16384), the size of the library's static "
               "scratch; use dla_voigt for longer grids");
  if (!c->noalloc_scratch) {
    c->err = "dla_voigt_noalloc: the static scratch has no device address on this context";
    return DLA_ERR_STATE;
  }
  if (num_lines <= 0 || num_lines > DLA_MAX_LINES) num_lines = DLA_MAX_LINES;  // voigt.c:266
  CU_TRY(c, cudaSetDevice(c->device));
  double* lam = c->noalloc_scratch;
  double* raw = lam + DLA_NOALLOC_MAX_POINTS;
  double* prof = raw + DLA_NOALLOC_MAX_POINTS;
  // nothing has to start at zero: the kernel writes raw[0, num_points) and prof[0, num_points - 6) before it reads them;
  // what an earlier call left beyond that is never touched
  int rc = DLA_OK;
  auto fail = [&](cudaError_t e, const char* what) {
    c->err = std::string("dla_voigt_noalloc: ") + what + ": " + cudaGetErrorString(e);
    rc = DLA_ERR_CUDA;
  };
  cudaError_t e;
  if ((e = cudaMemcpyAsync(lam, lambdas, (size_t)num_points * 8, cudaMemcpyDefault, c->stream)) != cudaSuccess) fail(e, "copy in");
  if (rc == DLA_OK) {
    dla::voigt_single_kernel<<<1, 1024, 0, c->stream>>>(lam, (int)num_points, z, N, num_lines, raw, prof);
    if ((e = cudaGetLastError()) != cudaSuccess) fail(e, "launch");
  }
  if (rc == DLA_OK && (e = cudaMemcpyAsync(profile, prof, (size_t)(num_points - 6) * 8, cudaMemcpyDefault, c->stream)) != cudaSuccess)
    fail(e, "copy out");
  if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess && rc == DLA_OK) fail(e, "sync");
  return rc;
}

static int dla_log_mvnpdf_low_rank_impl(dla_ctx* c, const double* y, const double* mu, const double* M, const double* d,
                            int64_t n, int32_t k, double* log_p) {
  if (!c) return DLA_ERR_BAD_ARG;
  if (!y || !mu || !M || !d || !log_p) BAD_ARG(c, "dla_log_mvnpdf_low_rank: NULL pointer");
  if (n < 1 || n > (1 << 24)) BAD_ARG(c, "dla_log_mvnpdf_low_rank: n out of range");
  if (k < 1 || k > dla::KGEN_MAX) BAD_ARG(c, "dla_log_mvnpdf_low_rank: k must be in 1..64");
  CU_TRY(c, cudaSetDevice(c->device));
  c->staged = false;  // reuses the chunk workspace
  c->ran = false;
  // One quasar, zero DLA samples: the no-DLA pseudo-sample (a == 1) with omega2 = 0 and
  // noise_variance = d reproduces log N(y; mu, M M' + diag(d)) exactly.
  QuasarDesc hd{};
  hd.n = (int32_t)n;
  hd.P = (int32_t)n + 6;
  hd.npad = (int32_t)round_up(n, dla::KC);
  hd.ident = 1;
  Chunk ch{0, 1, (int64_t)dla::TS * hd.npad, (int64_t)hd.npad * dla::NCOL, hd.npad, hd.P, false};
  int rc = ensure_workspace(c, k, ch.wu_doubles, ch.bmat_doubles, 1, 0, true);
  if (rc != DLA_OK) return rc;
  DevBuf &dd = c->seam[0], &dy = c->seam[1], &dmu = c->seam[2], &dM = c->seam[3], &dzero = c->seam[4], &dnv = c->seam[5],
         &dout = c->seam[6], &dpix = c->seam[7];
  auto cleanup = [&]() { cudaStreamSynchronize(c->stream); };  // nothing may still read the caller's arrays
#define SEAM_TRY(call)                                                                     \
  do {                                                                                     \
    cudaError_t e_ = (call);                                                               \
    if (e_ != cudaSuccess) {                                                               \
      c->err = std::string("dla_log_mvnpdf_low_rank: " #call ": ") + cudaGetErrorString(e_); \
      cleanup();                                                                           \
      return e_ == cudaErrorMemoryAllocation ? DLA_ERR_OOM : DLA_ERR_CUDA;                 \
    }                                                                                      \
  } while (0)
  SEAM_TRY(dd.ensure(sizeof(QuasarDesc)));
  SEAM_TRY(dy.ensure((size_t)n * 8));
  SEAM_TRY(dmu.ensure((size_t)n * 8));
  SEAM_TRY(dM.ensure((size_t)n * k * 8));
  SEAM_TRY(dzero.ensure((size_t)n * 8));
  SEAM_TRY(dnv.ensure((size_t)n * 8));
  SEAM_TRY(dout.ensure(64));
  SEAM_TRY(dpix.ensure((size_t)n * 32));
  SEAM_TRY(cudaMemcpyAsync(dd.p, &hd, sizeof hd, cudaMemcpyHostToDevice, c->stream));
  SEAM_TRY(cudaMemcpyAsync(dy.p, y, (size_t)n * 8, cudaMemcpyHostToDevice, c->stream));
  SEAM_TRY(cudaMemcpyAsync(dmu.p, mu, (size_t)n * 8, cudaMemcpyHostToDevice, c->stream));
  SEAM_TRY(cudaMemcpyAsync(dM.p, M, (size_t)n * k * 8, cudaMemcpyHostToDevice, c->stream));
  SEAM_TRY(cudaMemsetAsync(dzero.p, 0, (size_t)n * 8, c->stream));
  SEAM_TRY(cudaMemcpyAsync(dnv.p, d, (size_t)n * 8, cudaMemcpyHostToDevice, c->stream));
  Inputs in{dd.as<QuasarDesc>(), dzero.as<double>(), dy.as<double>(), dmu.as<double>(), dzero.as<double>(),
            dnv.as<double>(), dM.as<double>(), reinterpret_cast<const int32_t*>(dzero.p), dpix.as<double2>()};
  rc = launch_chunk(c, in, ch, 0, k, 1, dzero.as<double>(), dzero.as<double>(),
                    reinterpret_cast<const int32_t*>(dzero.p), c->d_ll_chunk.as<double>(),
                    dout.as<double>(), nullptr, nullptr, nullptr, nullptr, nullptr, nullptr);
  if (rc != DLA_OK) {
    cleanup();
    return rc;
  }
  SEAM_TRY(cudaMemcpyAsync(log_p, dout.p, 8, cudaMemcpyDeviceToHost, c->stream));
  SEAM_TRY(cudaStreamSynchronize(c->stream));
#undef SEAM_TRY
  return DLA_OK;
}


}  // extern "C" (implementations)

// No C++ exception crosses the ABI (std::bad_alloc of the host-side vectors, std::system_error of the helper thread):
// every entry point that can throw runs inside this guard.  After any failure the streams are drained, so that no
// copy still reads or writes caller memory when the call returns (the header's promise).
static void drain_after_failure(dla_ctx* c, bool reset_state) {
  if (!c) return;
  for (cudaStream_t st : {c->s_in, c->stream, c->stream2, c->s_out})
    if (st) cudaStreamSynchronize(st);
  cudaGetLastError();
  for (auto& set : c->sset) set.busy = false;
  c->slice_chunks = 0;
  c->hin = {};
  if (reset_state) {   // a CUDA / allocation failure: whatever was staged or computed is not to be trusted
    c->staged = false;
    c->ran = false;
  }
}
template <typename F>
static int guarded(dla_ctx* c, F&& f) {
  int rc;
  try {
    rc = f();
  } catch (const std::bad_alloc&) {
    if (c) c->err = "host allocation failed (std::bad_alloc)";
    rc = DLA_ERR_OOM;
  } catch (const std::exception& e) {
    if (c) c->err = std::string("unexpected C++ exception: ") + e.what();
    rc = DLA_ERR_CUDA;
  } catch (...) {
    if (c) c->err = "unexpected C++ exception";
    rc = DLA_ERR_CUDA;
  }
  if (rc != DLA_OK) drain_after_failure(c, rc != DLA_ERR_BAD_ARG && rc != DLA_ERR_STATE);
  return rc;
}

extern "C" {

int dla_set_samples(dla_ctx* c, const double* offset_samples, const double* nhi_samples) {
  return guarded(c, [&] { return dla_set_samples_impl(c, offset_samples, nhi_samples); });
}

int dla_stage_batch(dla_ctx* c, int64_t Q, const int64_t* pix_ptr, const int64_t* pad_ptr, const double* padded_wavelengths, const int32_t* profile_index, const double* flux, const double* mu, const double* omega2, const double* noise_variance, const double* M, const double* min_z_dla, const double* max_z_dla, const double* log_prior_dla, const double* log_prior_no_dla) {
  return guarded(c, [&] { return dla_stage_batch_impl(c, Q, pix_ptr, pad_ptr, padded_wavelengths, profile_index, flux, mu, omega2, noise_variance, M, min_z_dla, max_z_dla, log_prior_dla, log_prior_no_dla); });
}

int dla_run_staged(dla_ctx* c) {
  return guarded(c, [&] { return dla_run_staged_impl(c); });
}

int dla_fetch_staged(dla_ctx* c, double* sample_ll, double* ll_no_dla, double* ll_dla, double* log_post_no_dla, double* log_post_dla, double* model_posteriors, int32_t* map_index) {
  return guarded(c, [&] { return dla_fetch_staged_impl(c, sample_ll, ll_no_dla, ll_dla, log_post_no_dla, log_post_dla, model_posteriors, map_index); });
}

int dla_eval_batch(dla_ctx* c, int64_t Q, const int64_t* pix_ptr, const int64_t* pad_ptr, const double* padded_wavelengths, const int32_t* profile_index, const double* flux, const double* mu, const double* omega2, const double* noise_variance, const double* M, const double* min_z_dla, const double* max_z_dla, const double* log_prior_dla, const double* log_prior_no_dla, double* sample_ll, double* ll_no_dla, double* ll_dla, double* log_post_no_dla, double* log_post_dla, double* model_posteriors, int32_t* map_index) {
  return guarded(c, [&] { return dla_eval_batch_impl(c, Q, pix_ptr, pad_ptr, padded_wavelengths, profile_index, flux, mu, omega2, noise_variance, M, min_z_dla, max_z_dla, log_prior_dla, log_prior_no_dla, sample_ll, ll_no_dla, ll_dla, log_post_no_dla, log_post_dla, model_posteriors, map_index); });
}

int dla_set_model(dla_ctx* c, int64_t n_rest, const double* rest_wavelengths, const double* mu, const double* M, const double* log_omega, const dla_prep_params* prep) {
  return guarded(c, [&] { return dla_set_model_impl(c, n_rest, rest_wavelengths, mu, M, log_omega, prep); });
}

int dla_eval_spectra(dla_ctx* c, int64_t Q, const int64_t* spec_ptr, const double* wavelengths, const double* flux, const double* noise_variance, const uint8_t* pixel_mask, const double* z_qso, const double* log_prior_dla, const double* log_prior_no_dla, int32_t mask_gather, double* sample_ll, double* ll_no_dla, double* ll_dla, double* log_post_no_dla, double* log_post_dla, double* model_posteriors, int32_t* map_index, double* min_z_out, double* max_z_out) {
  return guarded(c, [&] { return dla_eval_spectra_impl(c, Q, spec_ptr, wavelengths, flux, noise_variance, pixel_mask, z_qso, log_prior_dla, log_prior_no_dla, mask_gather, sample_ll, ll_no_dla, ll_dla, log_post_no_dla, log_post_dla, model_posteriors, map_index, min_z_out, max_z_out); });
}

int dla_train_set_data(dla_ctx* c, int64_t Q, int64_t n, const double* y, const double* l1pz, const double* nv) {
  return guarded(c, [&] { return dla_train_set_data_impl(c, Q, n, y, l1pz, nv); });
}

int dla_train_objective(dla_ctx* c, const double* x, int32_t apply_priors, double* f, double* g) {
  return guarded(c, [&] { return dla_train_objective_impl(c, x, apply_priors, f, g); });
}

int dla_voigt(dla_ctx* c, const double* lambdas, int64_t num_points, double z, double N, int32_t num_lines, double* profile) {
  return guarded(c, [&] { return dla_voigt_impl(c, lambdas, num_points, z, N, num_lines, profile); });
}

int dla_voigt_noalloc(dla_ctx* c, const double* lambdas, int64_t num_points, double z, double N, int32_t num_lines, double* profile) {
  return guarded(c, [&] { return dla_voigt_noalloc_impl(c, lambdas, num_points, z, N, num_lines, profile); });
}

int dla_log_mvnpdf_low_rank(dla_ctx* c, const double* y, const double* mu, const double* M, const double* d, int64_t n, int32_t k, double* log_p) {
  return guarded(c, [&] { return dla_log_mvnpdf_low_rank_impl(c, y, mu, M, d, n, k, log_p); });
}

}  // extern "C"
