/* gp_dla_detection_b200/mex/voigt.c -- MEX gateway with the EXACT signature of the reference's
 * voigt MEX (voigt.c:253-304):
 *
 *     profile = voigt(lambdas, z, N [, num_lines])
 *
 * but evaluated on the B200 through dla_voigt() of libdla_b200.so.  With this file on the
 * MATLAB/Octave path instead of the reference's voigt.c, the UNMODIFIED process_qsos.m runs its
 * Voigt profiles on the GPU (seam-level parity; the fast path is dla_eval_mex.c).
 *
 * Build (MATLAB):  mex voigt.c -I../../include -L.. -ldla_b200
 * Build (Octave):  mkoctfile --mex voigt.c -I../../include -L.. -ldla_b200
 */
#include "mex.h"
#include "dla_b200.h"

static dla_ctx *g_ctx = 0;

static void cleanup(void) {
  if (g_ctx) dla_destroy(g_ctx);
  g_ctx = 0;
}

void mexFunction(int nlhs, mxArray *plhs[], int nrhs, const mxArray *prhs[]) {
  double *lambdas, *profile, z, N;
  int num_lines, rc;
  mwSize num_points;
  (void)nlhs;
  if (nrhs < 3) mexErrMsgIdAndTxt("dla_b200:voigt", "usage: profile = voigt(lambdas, z, N [, num_lines])");
  if (!g_ctx) {
    dla_params p = {20, 3, 0, 0, 0, 0};
    if (dla_create(&g_ctx, 0, &p) != DLA_OK)
      mexErrMsgIdAndTxt("dla_b200:create", "%s", dla_last_error(0));
    mexAtExit(cleanup);
  }
  lambdas = mxGetPr(prhs[0]);                                   /* voigt.c:262 */
  z = mxGetScalar(prhs[1]);                                     /* voigt.c:263 */
  N = mxGetScalar(prhs[2]);                                     /* voigt.c:264 */
  num_lines = (nrhs > 3) ? (int)mxGetScalar(prhs[3]) : 31;      /* voigt.c:16, :266 */
  num_points = mxGetNumberOfElements(prhs[0]);                  /* voigt.c:268 */
  if (num_points < 7) mexErrMsgIdAndTxt("dla_b200:voigt", "need at least 7 wavelengths");
  plhs[0] = mxCreateDoubleMatrix(num_points - 6, 1, mxREAL);    /* voigt.c:271 */
  profile = mxGetPr(plhs[0]);
  rc = dla_voigt(g_ctx, lambdas, (int64_t)num_points, z, N, num_lines, profile);
  if (rc != DLA_OK) mexErrMsgIdAndTxt("dla_b200:voigt", "%s", dla_last_error(g_ctx));
}
