/* gp_dla_detection_b200/mex/dla_eval_mex.c -- one-call replacement for the sample loop of
 * process_qsos.m (:184-198) and its reductions (:150-152, :202-209; generate_ascii_catalog.m:73):
 *
 *   [sample_ll, ll_dla, map_ind, ll_no_dla] = dla_eval_mex(padded_wavelengths, profile_ind, ...
 *       this_flux, this_mu, this_M, this_omega2, this_noise_variance, ...
 *       min_z_dla, max_z_dla, offset_samples, nhi_samples, num_lines)
 *
 *   profile_ind  1-based positions of the kept pixels in the n_full-long absorption profile:
 *                find(ind) with the `ind` of process_qsos.m:180 (as written: 1:n)
 *   this_M       n x k, passed through without a copy (MATLAB memory is column-major)
 *   sample_ll    1 x S  = sample_log_likelihoods_dla(quasar_ind, :)
 *   ll_dla       = log_likelihoods_dla(quasar_ind); map_ind 1-based; ll_no_dla = :150-152
 *
 * Build:  mex dla_eval_mex.c -I../../include -L.. -ldla_b200      (mkoctfile --mex ... for Octave)
 */
#include <stdlib.h>
#include "mex.h"
#include "dla_b200.h"

static dla_ctx *g_ctx = 0;
static int64_t g_S = -1;
static int g_k = -1, g_lines = -1;

static void cleanup(void) {
  if (g_ctx) dla_destroy(g_ctx);
  g_ctx = 0;
}

void mexFunction(int nlhs, mxArray *plhs[], int nrhs, const mxArray *prhs[]) {
  int64_t n, P, S, pix_ptr[2], pad_ptr[2], i;
  int k, num_lines, rc;
  int32_t *pidx, map_index = -1;
  double zero = 0.0, ll_no = 0, ll_dla = 0, lp_no = 0, lp_dla = 0, post[2];
  double min_z, max_z;
  if (nrhs != 12) mexErrMsgIdAndTxt("dla_b200:eval", "dla_eval_mex takes 12 arguments (see the file header)");
  P = (int64_t)mxGetNumberOfElements(prhs[0]);
  n = (int64_t)mxGetNumberOfElements(prhs[2]);
  k = (int)mxGetN(prhs[4]);
  S = (int64_t)mxGetNumberOfElements(prhs[9]);
  num_lines = (int)mxGetScalar(prhs[11]);
  if ((int64_t)mxGetM(prhs[4]) != n || (int64_t)mxGetNumberOfElements(prhs[1]) != n ||
      (int64_t)mxGetNumberOfElements(prhs[10]) != S)
    mexErrMsgIdAndTxt("dla_b200:eval", "inconsistent argument sizes");
  if (!g_ctx || g_S != S || g_k != k || g_lines != num_lines) {
    dla_params p = {0, 0, 0, 1, 0, 0};
    cleanup();
    p.k = k; p.num_lines = num_lines; p.num_samples = S;
    if (dla_create(&g_ctx, 0, &p) != DLA_OK) mexErrMsgIdAndTxt("dla_b200:create", "%s", dla_last_error(0));
    g_S = S; g_k = k; g_lines = num_lines;
    mexAtExit(cleanup);
  }
  /* the hypothesis set (process_qsos.m:37-40); cheap enough to resend per quasar */
  if (dla_set_samples(g_ctx, mxGetPr(prhs[9]), mxGetPr(prhs[10])) != DLA_OK)
    mexErrMsgIdAndTxt("dla_b200:samples", "%s", dla_last_error(g_ctx));
  pidx = (int32_t *)malloc((size_t)(n > 0 ? n : 1) * sizeof(int32_t));
  for (i = 0; i < n; i++) pidx[i] = (int32_t)(mxGetPr(prhs[1])[i]) - 1; /* 1-based -> 0-based */
  pix_ptr[0] = 0; pix_ptr[1] = n; pad_ptr[0] = 0; pad_ptr[1] = P;
  min_z = mxGetScalar(prhs[7]); max_z = mxGetScalar(prhs[8]);
  plhs[0] = mxCreateDoubleMatrix(1, (mwSize)S, mxREAL);
  rc = dla_eval_batch(g_ctx, 1, pix_ptr, pad_ptr, mxGetPr(prhs[0]), pidx, mxGetPr(prhs[2]), mxGetPr(prhs[3]),
                      mxGetPr(prhs[5]), mxGetPr(prhs[6]), mxGetPr(prhs[4]), &min_z, &max_z, &zero, &zero,
                      mxGetPr(plhs[0]), &ll_no, &ll_dla, &lp_no, &lp_dla, post, &map_index);
  free(pidx);
  if (rc != DLA_OK) mexErrMsgIdAndTxt("dla_b200:eval", "%s", dla_last_error(g_ctx));
  if (nlhs > 1) plhs[1] = mxCreateDoubleScalar(ll_dla);
  if (nlhs > 2) plhs[2] = mxCreateDoubleScalar((double)(map_index + 1));
  if (nlhs > 3) plhs[3] = mxCreateDoubleScalar(ll_no);
}
