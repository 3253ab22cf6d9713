/* include/dla_b200.h -- C ABI of libdla_b200.so
 *
 * B200 (sm_100a) drop-in for the per-spectrum DLA model-evaluation hot path of
 * sbird/gp_dla_detection.  Every entry point cites the reference interface it
 * replaces (paths relative to the reference tree).  Plain pointers and sizes
 * only; all pointers are HOST memory owned by the caller unless stated; the
 * library neither frees nor retains caller memory after a call returns.
 *
 * Conventions: functions return DLA_OK (0) or a negative dla_status; the text
 * of the last error of a context is available from dla_last_error().  No C++
 * exception crosses the ABI.  A dla_ctx is bound to one CUDA device and is not
 * thread-safe; different contexts may be driven from different host threads.
 * There is NO CPU fallback: without a usable CUDA device dla_create fails.
 */
#ifndef DLA_B200_H
#define DLA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLA_B200_ABI_VERSION 2

typedef enum {
  DLA_OK = 0,
  DLA_ERR_BAD_ARG = -1,   /* NULL pointer, non-monotone CSR, k/num_lines out of range, ... */
  DLA_ERR_CUDA = -2,      /* a CUDA runtime call or kernel failed */
  DLA_ERR_NO_DEVICE = -3, /* no CUDA device / wrong architecture (needs sm_100) */
  DLA_ERR_OOM = -4,       /* device or pinned-host allocation failed */
  DLA_ERR_STATE = -5      /* call order violated (e.g. samples not set, nothing staged) */
} dla_status;

typedef struct dla_ctx dla_ctx;

/* Run-constant parameters: set_parameters.m:36 (k), :48 (num_dla_samples), :63 (num_lines). */
typedef struct {
  int32_t k;               /* rank of the null model, 1..20 on the fused path */
  int32_t num_lines;       /* Lyman-series members, 1..31 (voigt.c:16); pipeline uses 3 */
  int64_t num_samples;     /* S = num_dla_samples */
  int32_t store_sample_ll; /* 1: return sample_log_likelihoods_dla; 0: only fused LSE/MAP outputs */
  int32_t reserved;
  int64_t workspace_bytes; /* device scratch budget for the W/U weight tiles; 0 = default (16 GiB) */
} dla_params;

/* Device time of the last dla_eval_batch / dla_run_staged, measured with CUDA
 * events recorded on the library's own stream (ms summed over chunks). */
typedef struct {
  double ms_total;       /* first launch -> last launch complete (compute only for run_staged) */
  double ms_phi;         /* K0 outer-product table                        */
  double ms_voigt;       /* K1 fused Voigt + weights                      */
  double ms_contract;    /* K2 DMMA contraction + Cholesky epilogue       */
  double ms_reduce;      /* K3 log-mean-exp + argmax + posteriors         */
  double ms_h2d, ms_d2h; /* copies (eval_batch only)                      */
  int64_t launches;      /* kernels launched                              */
  int64_t h2d_bytes, d2h_bytes;
  int64_t chunks;        /* quasar chunks the batch was split into        */
  double ms_chol;        /* K2c Cholesky rows (int8 slice path; 0 on the DMMA path, where it is K2's epilogue) */
  int64_t contraction;   /* 0: FP64 DMMA kernels, 1: exact-slice int8 tcgen05 contraction */
} dla_timing;

int dla_abi_version(void);
int dla_device_count(void);

/* Create a context on CUDA device `device`.  Replaces nothing in the reference
 * (it has no state); exists because device buffers and streams are reused. */
int dla_create(dla_ctx **out, int device, const dla_params *p);
void dla_destroy(dla_ctx *ctx);
/* Pointer to static/ctx-owned storage; never freed by the caller.  ctx may be
 * NULL to read the error of a failed dla_create. */
const char *dla_last_error(const dla_ctx *ctx);

/* The shared hypothesis set, length S each: offset_samples and nhi_samples of
 * dla_samples.mat (generate_dla_samples.m:13, :57; loaded process_qsos.m:37-40). */
int dla_set_samples(dla_ctx *ctx, const double *offset_samples, const double *nhi_samples);

/* Evaluate Q quasars (ragged, CSR-packed).  Replaces, per quasar,
 * process_qsos.m:150-155 (no-DLA likelihood), :163-165 (sample redshifts),
 * :184-198 (parfor: voigt MEX + absorbed model + log_mvnpdf_low_rank),
 * :202-212 (log-mean-exp, posteriors), :223-229 (model posteriors) and
 * generate_ascii_catalog.m:73 (MAP sample index).
 *
 *  pix_ptr   [Q+1]  offsets into the per-kept-pixel arrays; n_q = pix_ptr[q+1]-pix_ptr[q]
 *  pad_ptr   [Q+1]  offsets into padded_wavelengths; P_q = n_full_q + 6
 *  padded_wavelengths [sum P_q]   process_qsos.m:169-177 (Angstrom)
 *  profile_index [sum n_q]  0-based position of each kept pixel in the n_full_q
 *                           profile (process_qsos.m:180, :189); strictly increasing,
 *                           < n_full_q
 *  flux, mu, omega2, noise_variance [sum n_q]   this_flux, this_mu, this_omega2,
 *                           this_noise_variance (process_qsos.m:115-116, :139, :147)
 *  M         per quasar n_q x k COLUMN-major (MATLAB memory), quasar q at k*pix_ptr[q]
 *  min_z_dla, max_z_dla [Q]             process_qsos.m:160-161
 *  log_prior_dla, log_prior_no_dla [Q]  process_qsos.m:129-132
 * Outputs (caller-allocated; any may be NULL to skip it):
 *  sample_ll [Q*S]  quasar-major rows (process_qsos.m:195); needs store_sample_ll=1
 *  ll_no_dla, ll_dla [Q]               :150, :208
 *  log_post_no_dla, log_post_dla [Q]   :154, :211
 *  model_posteriors [Q*2] (no_dla, dla) :226-229
 *  map_index [Q]    0-based first argmax over samples, NaN ignored (-1 if all NaN)
 */
int dla_eval_batch(dla_ctx *ctx, int64_t Q,
                   const int64_t *pix_ptr, const int64_t *pad_ptr,
                   const double *padded_wavelengths, const int32_t *profile_index,
                   const double *flux, const double *mu, const double *omega2,
                   const double *noise_variance, const double *M,
                   const double *min_z_dla, const double *max_z_dla,
                   const double *log_prior_dla, const double *log_prior_no_dla,
                   double *sample_ll, double *ll_no_dla, double *ll_dla,
                   double *log_post_no_dla, double *log_post_dla,
                   double *model_posteriors, int32_t *map_index);

/* The same computation split so that inputs can stay resident in HBM:
 * dla_stage_batch copies the inputs to the device (same arguments as
 * dla_eval_batch's inputs), dla_run_staged runs the kernels on the resident
 * data (may be called repeatedly), dla_fetch_staged copies results back. */
int dla_stage_batch(dla_ctx *ctx, int64_t Q,
                    const int64_t *pix_ptr, const int64_t *pad_ptr,
                    const double *padded_wavelengths, const int32_t *profile_index,
                    const double *flux, const double *mu, const double *omega2,
                    const double *noise_variance, const double *M,
                    const double *min_z_dla, const double *max_z_dla,
                    const double *log_prior_dla, const double *log_prior_no_dla);
int dla_run_staged(dla_ctx *ctx);
int dla_fetch_staged(dla_ctx *ctx, double *sample_ll, double *ll_no_dla, double *ll_dla,
                     double *log_post_no_dla, double *log_post_dla,
                     double *model_posteriors, int32_t *map_index);

/* ---- per-quasar preparation on the device -------------------------------
 * dla_set_model + dla_eval_spectra replace process_qsos.m:102-120 (rest-frame cut, pixel
 * mask), :139-147 (linear interpolation of mu / M / log_omega onto the spectrum, omega^2
 * scaling), :160-161 (z_DLA range), :169-180 (padded grid, gather index) AND everything
 * dla_eval_batch replaces, from the raw ragged arrays of preloaded_qsos.mat
 * (preload_qsos.m:73-79).  Selection, counts and the padded grid are one cheap host pass
 * inside the library; the interpolation of the n x (k+2) model values, exp and scaling run
 * in a CUDA kernel, so the learned model is uploaded once and per quasar only the kept
 * wavelengths, flux and noise variance cross PCIe. */
typedef struct {
  double min_lambda, max_lambda;      /* set_parameters.m:33-34 */
  double lya_wavelength, lyman_limit; /* :5, :7 */
  double max_z_cut, min_z_cut;        /* :65, :69  kms_to_z(3000) */
  double pixel_spacing;               /* :60 */
  int32_t width;                      /* :59 (must be 3: voigt.c:229) */
  int32_t reserved;
  double log_c_0, log_tau_0, log_beta; /* learned_qso_model: process_qsos.m:84-86 */
} dla_prep_params;

/* learned_qso_model_*.mat (learn_qso_model.m:113-123): rest_wavelengths [n_rest] strictly
 * increasing, mu [n_rest], M [n_rest x k] COLUMN-major, log_omega [n_rest]. */
int dla_set_model(dla_ctx *ctx, int64_t n_rest, const double *rest_wavelengths, const double *mu,
                  const double *M, const double *log_omega, const dla_prep_params *prep);

/* Q raw spectra, ragged: spec_ptr [Q+1] offsets into wavelengths / flux / noise_variance /
 * pixel_mask (1 = masked), z_qso [Q].  mask_gather: 0 = the gather as written in the
 * reference (process_qsos.m:180/:189: first n profile elements), 1 = the unmasked positions.
 * Outputs as dla_eval_batch, plus min_z_dla / max_z_dla [Q] (process_qsos.m:160-161; may be
 * NULL).  A quasar with no unmasked pixel in the modelled range is DLA_ERR_BAD_ARG.
 * Q is unbounded: more than 2048 quasars are evaluated as consecutive sub-batches inside the
 * call (one short-lived helper host thread selects the next sub-batch while the device works
 * on the current one); the call is still synchronous, and on an error outputs of earlier
 * sub-batches may already have been written. */
int dla_eval_spectra(dla_ctx *ctx, int64_t Q, const int64_t *spec_ptr, const double *wavelengths,
                     const double *flux, const double *noise_variance, const uint8_t *pixel_mask,
                     const double *z_qso, const double *log_prior_dla, const double *log_prior_no_dla,
                     int32_t mask_gather,
                     double *sample_ll, double *ll_no_dla, double *ll_dla,
                     double *log_post_no_dla, double *log_post_dla,
                     double *model_posteriors, int32_t *map_index,
                     double *min_z_dla, double *max_z_dla);

/* ---- training objective -------------------------------------------------
 * objective.m:12-75 over spectrum_loss.m:14-76: f(x) = -sum_i log N(y_i; 0, M M' + diag(d_i)) and
 * its gradient, x = [vec M (num_pixels x k, column-major); log_omega (num_pixels); log c_0;
 * log tau_0; log beta], k = the context's rank (<= 32 here).  The training set stays resident on
 * the device between calls (an L-BFGS run evaluates the objective many times):
 *   centered_rest_fluxes, lya_1pzs, rest_noise_variances  [num_quasars][num_pixels], ROW-major
 *   (one spectrum contiguous; MATLAB's arrays are the transpose), NaN in the flux = missing pixel
 *   (objective.m:42).  apply_priors != 0 adds the tau_0 / beta priors of objective.m:60-71.
 * A multi-GPU run gives every rank its own slice of the quasars and SUMS f and g over ranks (the
 * one all-reduce of the pipeline; priors added once) -- see gp_dla_detection_b200/learn_qso_model.py. */
int dla_train_set_data(dla_ctx *ctx, int64_t num_quasars, int64_t num_pixels,
                       const double *centered_rest_fluxes, const double *lya_1pzs,
                       const double *rest_noise_variances);
int dla_train_objective(dla_ctx *ctx, const double *x, int32_t apply_priors, double *f, double *g);

int dla_get_timing(const dla_ctx *ctx, dla_timing *out);

/* Unit seams (one call = one reference call; for parity tests and for running
 * the UNMODIFIED reference loop on the GPU through the MEX shims).
 *
 * dla_voigt = the voigt MEX, voigt.c:253-304:
 *   profile = voigt(lambdas, z, N, num_lines); profile has num_points-6 entries. */
int dla_voigt(dla_ctx *ctx, const double *lambdas, int64_t num_points, double z, double N,
              int32_t num_lines, double *profile);
/* dla_voigt_noalloc: NO COUNTERPART IN THE REFERENCE; added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same
 * arguments, kernel and result as dla_voigt, but the call allocates nothing (no cudaMalloc / cudaFree, no allocating
 * library routine): the staged wavelengths, the kernel's scratch (the raw transmissions of voigt.c:282-292) and the
 * profile live in ONE file-scope `static __device__` array of the library, 3 x DLA_NOALLOC_MAX_POINTS doubles =
 * 384 KB, which every GPU that loads the library holds until the process exits; its device address is looked up
 * once per device in dla_create.  `lambdas` and `profile` may be host or device pointers (cudaMemcpyDefault).
 * num_points > DLA_NOALLOC_MAX_POINTS is rejected with DLA_ERR_BAD_ARG.  Every call on a device shares the array,
 * which keeps the last call's contents (nothing in it has to start at zero: the kernel writes what it reads); calls
 * that overlap on the GPU -- two contexts of one device used from two threads -- RACE on it. */
#define DLA_NOALLOC_MAX_POINTS 16384
int dla_voigt_noalloc(dla_ctx *ctx, const double *lambdas, int64_t num_points, double z, double N,
                      int32_t num_lines, double *profile);
/* dla_log_mvnpdf_low_rank = log_mvnpdf_low_rank.m:5-34:
 *   log N(y; mu, M M' + diag(d)); M column-major n x k, k <= 20. */
int dla_log_mvnpdf_low_rank(dla_ctx *ctx, const double *y, const double *mu, const double *M,
                            const double *d, int64_t n, int32_t k, double *log_p);

#ifdef __cplusplus
}
#endif
#endif /* DLA_B200_H */
