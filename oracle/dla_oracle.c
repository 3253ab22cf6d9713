/* oracle/dla_oracle.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C CPU restatement of the reference hot path, line by line and AS
 * WRITTEN (no Woodbury recast, no symmetry tricks):
 *   oracle_voigt               <- voigt.c:253-304 (loops :278-299)
 *   oracle_log_mvnpdf_low_rank <- log_mvnpdf_low_rank.m:5-34
 *   oracle_process_quasar      <- process_qsos.m:150-155, :163-165, :184-212, :223-232
 *                                 and generate_ascii_catalog.m:73 (MAP index)
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library; the product (libdla_b200.so) never
 * does.  Compiled with -ffp-contract=off so the velocity subtraction
 * (voigt.c:287) and the 7-tap accumulation (voigt.c:299) round as in an x86-64
 * `mex` build without FMA.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "voigt_constants.h"

double voigt(double x, double sigma, double gamma); /* oracle/cerf_port.c */

typedef long (*voigt_fn_t)(const double *lambdas, long num_points, double z, double N,
                           int num_lines, double *profile_out);

/* voigt.c:253-304.  profile has num_points - 6 entries. */
long oracle_voigt(const double *lambdas, long num_points, double z, double N,
                  int num_lines, double *profile) {
  const double c = ORC_C_CGS, sigma = ORC_SIGMA;
  const int width = 3;
  long i, n_out = num_points - 2 * width;
  int j, k;
  double multipliers[ORC_MAX_LINES];
  double *raw = (double *)malloc((size_t)num_points * sizeof(double));
  if (num_lines <= 0 || num_lines > ORC_MAX_LINES) num_lines = ORC_MAX_LINES; /* voigt.c:16,:266 */
  for (j = 0; j < num_lines; j++) /* voigt.c:278-279 */
    multipliers[j] = c / (orc_transition_wavelengths[j] * (1 + z)) / 1e8;
  for (i = 0; i < num_points; i++) { /* voigt.c:282-292 */
    double total = 0;
    for (j = 0; j < num_lines; j++) {
      double velocity = lambdas[i] * multipliers[j] - c;
      total += -orc_leading_constants[j] * voigt(velocity, sigma, orc_gammas[j]);
    }
    raw[i] = exp(N * total);
  }
  for (i = 0; i < n_out; i++) { /* voigt.c:297-299; output zero-initialised (:271) */
    double acc = 0.0;
    for (k = 0; k <= 2 * width; k++) acc += raw[i + k] * orc_instrument_profile[k];
    profile[i] = acc;
  }
  free(raw);
  return n_out;
}

/* log_mvnpdf_low_rank.m:5-34, as written.  M column-major n x k.
 * work: caller scratch of at least 2*n*k + 2*k*k + 4*n + 2*k doubles. */
double oracle_log_mvnpdf_low_rank_w(const double *y_in, const double *mu, const double *M,
                                    const double *d, long n, int k, double *work) {
  const double log_2pi = 1.83787706640934534; /* :7 */
  double *y = work;                 /* n   : y - mu            (:11) */
  double *D_inv_y = y + n;          /* n                        (:14) */
  double *D_inv_M = D_inv_y + n;    /* n*k col-major            (:15) */
  double *C = D_inv_M + n * k;      /* k*n col-major (k rows)   (:26) */
  double *B = C + n * k;            /* k*k col-major            (:22) */
  double *Cy = B + k * k;           /* k */
  double *K_inv_y = Cy + k;         /* n */
  long p;
  int i, j, m;
  double log_det_K, quad;

  for (p = 0; p < n; p++) {
    double d_inv = 1.0 / d[p]; /* :13 */
    y[p] = y_in[p] - mu[p];
    D_inv_y[p] = d_inv * y[p];
    for (j = 0; j < k; j++) D_inv_M[j * n + p] = d_inv * M[j * n + p];
  }
  /* B = M' * D_inv_M (:22), B(ii) += 1 (:23) */
  for (i = 0; i < k; i++)
    for (j = 0; j < k; j++) {
      double s = 0;
      for (p = 0; p < n; p++) s += M[i * n + p] * D_inv_M[j * n + p];
      B[j * k + i] = s;
    }
  for (i = 0; i < k; i++) B[i * k + i] += 1;
  /* L = chol(B): upper triangular, L'L = B (:24).  Stored in B's upper triangle. */
  for (j = 0; j < k; j++) {
    double s = B[j * k + j];
    for (m = 0; m < j; m++) s -= B[j * k + m] * B[j * k + m]; /* L(m,j)^2 */
    if (!(s > 0)) return NAN;                                 /* chol would error */
    s = sqrt(s);
    B[j * k + j] = s;
    for (i = j + 1; i < k; i++) { /* L(j,i) */
      double t = B[i * k + j];
      for (m = 0; m < j; m++) t -= B[j * k + m] * B[i * k + m];
      B[i * k + j] = t / s;
    }
  }
#define L_(r, c_) B[(c_) * k + (r)] /* upper: r <= c_ */
  /* C = L \ (L' \ D_inv_M')  (:26): k x n, column p solves B c = D_inv_M(p,:)' */
  for (p = 0; p < n; p++) {
    double *col = C + p * k;
    for (i = 0; i < k; i++) { /* forward: L' t = b */
      double t = D_inv_M[i * n + p];
      for (m = 0; m < i; m++) t -= L_(m, i) * col[m];
      col[i] = t / L_(i, i);
    }
    for (i = k - 1; i >= 0; i--) { /* backward: L c = t */
      double t = col[i];
      for (m = i + 1; m < k; m++) t -= L_(i, m) * col[m];
      col[i] = t / L_(i, i);
    }
  }
  /* K_inv_y = D_inv_y - D_inv_M * (C * y)  (:28) */
  for (i = 0; i < k; i++) Cy[i] = 0;
  for (p = 0; p < n; p++)
    for (i = 0; i < k; i++) Cy[i] += C[p * k + i] * y[p];
  for (p = 0; p < n; p++) {
    double s = 0;
    for (j = 0; j < k; j++) s += D_inv_M[j * n + p] * Cy[j];
    K_inv_y[p] = D_inv_y[p] - s;
  }
  /* log_det_K = sum(log(d)) + 2 * sum(log(diag(L)))  (:30) */
  log_det_K = 0;
  for (p = 0; p < n; p++) log_det_K += log(d[p]);
  {
    double s = 0;
    for (i = 0; i < k; i++) s += log(L_(i, i));
    log_det_K += 2 * s;
  }
#undef L_
  quad = 0;
  for (p = 0; p < n; p++) quad += y[p] * K_inv_y[p];
  return -0.5 * (quad + log_det_K + n * log_2pi); /* :32 */
}

static size_t lmlr_work_doubles(long n, int k) {
  return (size_t)(2 * n * k + 2 * k * k + 4 * n + 2 * k + 16);
}

double oracle_log_mvnpdf_low_rank(const double *y, const double *mu, const double *M,
                                  const double *d, long n, int k) {
  double *w = (double *)malloc(lmlr_work_doubles(n, k) * sizeof(double));
  double r = oracle_log_mvnpdf_low_rank_w(y, mu, M, d, n, k, w);
  free(w);
  return r;
}

/* One quasar of process_qsos.m:150-212 (+ :223-232 posterior, + MAP index).
 *   scalars_out[0] log_likelihoods_no_dla   (:150-152)
 *   scalars_out[1] log_likelihoods_dla      (:208-209)
 *   scalars_out[2] log_posteriors_no_dla    (:154-155)
 *   scalars_out[3] log_posteriors_dla       (:211-212)
 *   scalars_out[4..5] model_posteriors      (:223-229)
 * voigt_fn == NULL -> oracle_voigt above; else e.g. ref_voigt from
 * oracle/_ref/libvoigt_ref.so (the reference's voigt.c compiled verbatim).
 * first_sample/num_eval allow a bounded sample of the workload (bench cpu
 * baseline): only samples [first_sample, first_sample+num_eval) are computed
 * and the reductions run over those. Returns the 0-based MAP index. */
long oracle_process_quasar(long n, long n_full, const double *padded_wavelengths,
                           const int *profile_index, const double *flux, const double *mu,
                           const double *M, const double *omega2, const double *noise_variance,
                           double min_z_dla, double max_z_dla, long S, const double *offset_samples,
                           const double *nhi_samples, int num_lines, int k, double log_prior_dla,
                           double log_prior_no_dla, long first_sample, long num_eval,
                           voigt_fn_t voigt_fn, int nthreads, double *sample_ll /* [S] */,
                           double *scalars_out /* [6] */) {
  long P = n_full + 6, i;
  long map_index = -1;
  if (!voigt_fn) voigt_fn = oracle_voigt;
  if (num_eval <= 0) { first_sample = 0; num_eval = S; }
  (void)nthreads;
  /* process_qsos.m:150-152 */
  {
    double *dd = (double *)malloc((size_t)n * sizeof(double));
    for (i = 0; i < n; i++) dd[i] = omega2[i] + noise_variance[i];
    scalars_out[0] = oracle_log_mvnpdf_low_rank(flux, mu, M, dd, n, k);
    free(dd);
  }
#ifdef _OPENMP
#pragma omp parallel num_threads(nthreads > 0 ? nthreads : 1)
#endif
  {
    double *profile = (double *)malloc((size_t)(n_full > 0 ? n_full : 1) * sizeof(double));
    double *dla_mu = (double *)malloc((size_t)n * sizeof(double));
    double *dla_M = (double *)malloc((size_t)n * k * sizeof(double));
    double *dd = (double *)malloc((size_t)n * sizeof(double));
    double *work = (double *)malloc(lmlr_work_doubles(n, k) * sizeof(double));
    long s, p;
    int j;
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 8)
#endif
    for (s = first_sample; s < first_sample + num_eval; s++) { /* parfor, :184 */
      /* :163-165 */
      double z = min_z_dla + (max_z_dla - min_z_dla) * offset_samples[s];
      voigt_fn(padded_wavelengths, P, z, nhi_samples[s], num_lines, profile); /* :186-187 */
      for (p = 0; p < n; p++) {
        double a = profile[profile_index[p]]; /* :189 */
        dla_mu[p] = mu[p] * a;                /* :191 */
        for (j = 0; j < k; j++) dla_M[j * n + p] = M[j * n + p] * a; /* :192 */
        dd[p] = omega2[p] * (a * a) + noise_variance[p];            /* :193, :197 */
      }
      sample_ll[s] = oracle_log_mvnpdf_low_rank_w(flux, dla_mu, dla_M, dd, n, k, work); /* :195-197 */
    }
    free(profile); free(dla_mu); free(dla_M); free(dd); free(work);
  }
  /* :202-209; MATLAB max ignores NaN, exp(NaN - m) poisons mean */
  {
    double m = -INFINITY, acc = 0;
    int any = 0;
    for (i = first_sample; i < first_sample + num_eval; i++) {
      double v = sample_ll[i];
      if (v != v) continue;
      if (!any || v > m) { m = v; map_index = i; any = 1; } /* first maximum wins */
    }
    if (!any) m = NAN;
    for (i = first_sample; i < first_sample + num_eval; i++) acc += exp(sample_ll[i] - m);
    scalars_out[1] = m + log(acc / (double)num_eval);
  }
  scalars_out[2] = log_prior_no_dla + scalars_out[0]; /* :154-155 */
  scalars_out[3] = log_prior_dla + scalars_out[1];    /* :211-212 */
  { /* :223-229 */
    double a = scalars_out[2], b = scalars_out[3];
    double mx = (a > b || b != b) ? a : b; /* MATLAB max ignores NaN */
    double ea = exp(a - mx), eb = exp(b - mx);
    scalars_out[4] = ea / (ea + eb);
    scalars_out[5] = eb / (ea + eb);
  }
  return map_index;
}

int oracle_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
