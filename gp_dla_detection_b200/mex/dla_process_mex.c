/* gp_dla_detection_b200/mex/dla_process_mex.c -- the whole quasar loop of process_qsos.m (:88-220) and the model
 * posteriors (:223-229) in ONE call, at device speed: the cell arrays of preloaded_qsos.mat go in as they are, the
 * variables the script saves (:235-243) come out.  Runs dla_set_model + dla_eval_spectra, i.e. the per-quasar
 * preparation (:102-180) happens inside the library as well.
 *
 *   [min_z_dlas, max_z_dlas, log_likelihoods_no_dla, sample_log_likelihoods_dla, log_likelihoods_dla, ...
 *    log_posteriors_no_dla, log_posteriors_dla, model_posteriors, map_ind] = dla_process_mex( ...
 *       all_wavelengths, all_flux, all_noise_variance, all_pixel_mask, z_qsos, ...   % preloaded_qsos.mat (cells, vector)
 *       log_priors_dla, log_priors_no_dla, ...                                        % process_qsos.m:123-132
 *       rest_wavelengths, mu, M, log_omega, log_c_0, log_tau_0, log_beta, ...         % learned_qso_model_*.mat
 *       offset_samples, nhi_samples, ...                                              % dla_samples.mat
 *       params)
 *
 *   params   row vector [min_lambda max_lambda lya_wavelength lyman_limit max_z_cut min_z_cut pixel_spacing width
 *            num_lines mask_gather]: set_parameters.m:33-34, :5, :7, kms_to_z(3000) twice (:65, :69), :60, :59, :63;
 *            mask_gather 0 = the gather of process_qsos.m:180/:189 as written, 1 = the unmasked positions
 *   cells    num_quasars x 1, each a double vector (all_pixel_mask: logical or double vectors)
 *   sample_log_likelihoods_dla   num_quasars x num_dla_samples (MATLAB's column-major layout, as the script holds it)
 *   model_posteriors             num_quasars x 2 = [p_no_dlas, p_dlas] (:223-232); map_ind 1-based (generate_ascii_catalog.m:73)
 *
 * The context (device buffers, the sorted hypothesis set, the resident model) survives between calls and is rebuilt
 * only when k, num_lines or num_dla_samples change.
 * Build:  mex dla_process_mex.c -I../../include -L.. -ldla_b200      (mkoctfile --mex ... for Octave)
 */
#include <stdlib.h>
#include <string.h>
#include "mex.h"
#include "dla_b200.h"

static dla_ctx *g_ctx = 0;
static int64_t g_S = -1;
static int g_k = -1, g_lines = -1;

static void cleanup(void) {
  if (g_ctx) dla_destroy(g_ctx);
  g_ctx = 0;
}

static void fail(const char *what) { mexErrMsgIdAndTxt("dla_b200:process", "%s", what); }

void mexFunction(int nlhs, mxArray *plhs[], int nrhs, const mxArray *prhs[]) {
  int64_t Q, S, n_rest, total = 0, q, i, *spec_ptr;
  int k, num_lines, mask_gather, rc;
  const double *par;
  double *wl, *fl, *nv, *sample_rowmajor, *post;
  uint8_t *mask;
  int32_t *map_index;
  dla_prep_params prep;
  if (nrhs != 17) fail("dla_process_mex takes 17 arguments (see the file header)");
  Q = (int64_t)mxGetNumberOfElements(prhs[4]);
  if ((int64_t)mxGetNumberOfElements(prhs[0]) != Q || (int64_t)mxGetNumberOfElements(prhs[1]) != Q ||
      (int64_t)mxGetNumberOfElements(prhs[2]) != Q || (int64_t)mxGetNumberOfElements(prhs[3]) != Q ||
      (int64_t)mxGetNumberOfElements(prhs[5]) != Q || (int64_t)mxGetNumberOfElements(prhs[6]) != Q)
    fail("the four cell arrays, z_qsos and the two prior vectors must have num_quasars elements each");
  n_rest = (int64_t)mxGetNumberOfElements(prhs[7]);
  k = (int)mxGetN(prhs[9]);
  if ((int64_t)mxGetM(prhs[9]) != n_rest || (int64_t)mxGetNumberOfElements(prhs[8]) != n_rest ||
      (int64_t)mxGetNumberOfElements(prhs[10]) != n_rest)
    fail("mu, M and log_omega must have one row per rest wavelength");
  S = (int64_t)mxGetNumberOfElements(prhs[14]);
  if ((int64_t)mxGetNumberOfElements(prhs[15]) != S) fail("offset_samples and nhi_samples differ in length");
  if (mxGetNumberOfElements(prhs[16]) != 10) fail("params must have 10 elements (see the file header)");
  par = mxGetPr(prhs[16]);
  memset(&prep, 0, sizeof prep);
  prep.min_lambda = par[0]; prep.max_lambda = par[1]; prep.lya_wavelength = par[2]; prep.lyman_limit = par[3];
  prep.max_z_cut = par[4]; prep.min_z_cut = par[5]; prep.pixel_spacing = par[6]; prep.width = (int32_t)par[7];
  num_lines = (int)par[8];
  mask_gather = (int)par[9];
  prep.log_c_0 = mxGetScalar(prhs[11]); prep.log_tau_0 = mxGetScalar(prhs[12]); prep.log_beta = mxGetScalar(prhs[13]);

  if (!g_ctx || g_S != S || g_k != k || g_lines != num_lines) {
    dla_params p;
    memset(&p, 0, sizeof p);
    cleanup();
    p.k = k; p.num_lines = num_lines; p.num_samples = S; p.store_sample_ll = 1;
    if (dla_create(&g_ctx, 0, &p) != DLA_OK) mexErrMsgIdAndTxt("dla_b200:create", "%s", dla_last_error(0));
    g_S = S; g_k = k; g_lines = num_lines;
    mexAtExit(cleanup);
  }
  /* unchanged arrays are recognised inside the library: no re-sort, no copy */
  if (dla_set_samples(g_ctx, mxGetPr(prhs[14]), mxGetPr(prhs[15])) != DLA_OK) fail(dla_last_error(g_ctx));
  if (dla_set_model(g_ctx, n_rest, mxGetPr(prhs[7]), mxGetPr(prhs[8]), mxGetPr(prhs[9]), mxGetPr(prhs[10]), &prep) != DLA_OK)
    fail(dla_last_error(g_ctx));

  /* pack the ragged cell arrays (preload_qsos.m:73-79) */
  spec_ptr = (int64_t *)mxMalloc((size_t)(Q + 1) * sizeof(int64_t));
  spec_ptr[0] = 0;
  for (q = 0; q < Q; q++) {
    const mxArray *w = mxGetCell(prhs[0], (mwSize)q);
    const int64_t n = w ? (int64_t)mxGetNumberOfElements(w) : 0;
    const mxArray *f = mxGetCell(prhs[1], (mwSize)q), *v = mxGetCell(prhs[2], (mwSize)q), *m = mxGetCell(prhs[3], (mwSize)q);
    if (!f || !v || !m || (int64_t)mxGetNumberOfElements(f) != n || (int64_t)mxGetNumberOfElements(v) != n ||
        (int64_t)mxGetNumberOfElements(m) != n)
      fail("the four cell arrays must hold vectors of equal length per quasar");
    total += n;
    spec_ptr[q + 1] = total;
  }
  wl = (double *)mxMalloc((size_t)(total > 0 ? total : 1) * sizeof(double));
  fl = (double *)mxMalloc((size_t)(total > 0 ? total : 1) * sizeof(double));
  nv = (double *)mxMalloc((size_t)(total > 0 ? total : 1) * sizeof(double));
  mask = (uint8_t *)mxMalloc((size_t)(total > 0 ? total : 1));
  for (q = 0; q < Q; q++) {
    const int64_t a = spec_ptr[q], n = spec_ptr[q + 1] - a;
    const mxArray *m = mxGetCell(prhs[3], (mwSize)q);
    memcpy(wl + a, mxGetPr(mxGetCell(prhs[0], (mwSize)q)), (size_t)n * sizeof(double));
    memcpy(fl + a, mxGetPr(mxGetCell(prhs[1], (mwSize)q)), (size_t)n * sizeof(double));
    memcpy(nv + a, mxGetPr(mxGetCell(prhs[2], (mwSize)q)), (size_t)n * sizeof(double));
    if (mxIsLogical(m)) {
      const mxLogical *b = mxGetLogicals(m);
      for (i = 0; i < n; i++) mask[a + i] = b[i] ? 1 : 0;
    } else {
      const double *b = mxGetPr(m);
      for (i = 0; i < n; i++) mask[a + i] = b[i] != 0.0 ? 1 : 0;
    }
  }

  /* outputs: the script's variables (:74-82: NaN-initialised there; every entry is written here) */
  plhs[0] = mxCreateDoubleMatrix((mwSize)Q, 1, mxREAL);               /* min_z_dlas */
  { mxArray *t[9]; int o;
    t[0] = plhs[0];
    for (o = 1; o < 9; o++) t[o] = 0;
    t[1] = mxCreateDoubleMatrix((mwSize)Q, 1, mxREAL);                /* max_z_dlas */
    t[2] = mxCreateDoubleMatrix((mwSize)Q, 1, mxREAL);                /* log_likelihoods_no_dla */
    t[3] = mxCreateDoubleMatrix((mwSize)Q, (mwSize)S, mxREAL);        /* sample_log_likelihoods_dla */
    t[4] = mxCreateDoubleMatrix((mwSize)Q, 1, mxREAL);                /* log_likelihoods_dla */
    t[5] = mxCreateDoubleMatrix((mwSize)Q, 1, mxREAL);                /* log_posteriors_no_dla */
    t[6] = mxCreateDoubleMatrix((mwSize)Q, 1, mxREAL);                /* log_posteriors_dla */
    t[7] = mxCreateDoubleMatrix((mwSize)Q, 2, mxREAL);                /* model_posteriors */
    t[8] = mxCreateDoubleMatrix((mwSize)Q, 1, mxREAL);                /* map_ind */
    sample_rowmajor = (double *)mxMalloc((size_t)(Q * S > 0 ? Q * S : 1) * sizeof(double));
    post = (double *)mxMalloc((size_t)(Q > 0 ? 2 * Q : 1) * sizeof(double));
    map_index = (int32_t *)mxMalloc((size_t)(Q > 0 ? Q : 1) * sizeof(int32_t));
    rc = dla_eval_spectra(g_ctx, Q, spec_ptr, wl, fl, nv, mask, mxGetPr(prhs[4]), mxGetPr(prhs[5]), mxGetPr(prhs[6]),
                          mask_gather, sample_rowmajor, mxGetPr(t[2]), mxGetPr(t[4]), mxGetPr(t[5]), mxGetPr(t[6]), post,
                          map_index, mxGetPr(t[0]), mxGetPr(t[1]));
    if (rc == DLA_OK) {
      double *sl = mxGetPr(t[3]), *mp = mxGetPr(t[7]), *mi = mxGetPr(t[8]);
      for (q = 0; q < Q; q++) {                    /* quasar-major rows -> MATLAB's column-major num_quasars x S */
        for (i = 0; i < S; i++) sl[i * Q + q] = sample_rowmajor[q * S + i];
        mp[q] = post[2 * q];                       /* p_no_dlas (:231) */
        mp[Q + q] = post[2 * q + 1];               /* p_dlas    (:232) */
        mi[q] = (double)(map_index[q] + 1);
      }
    }
    mxFree(sample_rowmajor); mxFree(post); mxFree(map_index);
    mxFree(spec_ptr); mxFree(wl); mxFree(fl); mxFree(nv); mxFree(mask);
    if (rc != DLA_OK) fail(dla_last_error(g_ctx));
    for (o = 1; o < 9; o++) {
      if (o < nlhs) plhs[o] = t[o];
    }
  }
}
