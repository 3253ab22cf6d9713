/* oracle/shim/mex_driver.c -- TEST INFRASTRUCTURE ONLY.
 * Plain-C entry point that calls the reference's mexFunction (voigt.c:253)
 * the way MATLAB would for `profile = voigt(lambdas, z, N, num_lines)`
 * (process_qsos.m:186-187).  Linked together with /root/reference/voigt.c
 * (compiled verbatim) into oracle/_ref/libvoigt_ref.so by oracle/Makefile.
 */
#include "mex.h"
#include <string.h>

/* returns the number of profile points written (num_points - 6) */
long ref_voigt(const double *lambdas, long num_points, double z, double N,
               int num_lines, double *profile_out) {
  mxArray a_l, a_z, a_n, a_k;
  double zz = z, nn = N, kk = (double)num_lines;
  const mxArray *prhs[4];
  mxArray *plhs[1] = {0};
  a_l.pr = (double *)lambdas; a_l.m = (mwSize)num_points; a_l.n = 1;
  a_z.pr = &zz; a_z.m = a_z.n = 1;
  a_n.pr = &nn; a_n.m = a_n.n = 1;
  a_k.pr = &kk; a_k.m = a_k.n = 1;
  prhs[0] = &a_l; prhs[1] = &a_z; prhs[2] = &a_n; prhs[3] = &a_k;
  mexFunction(1, plhs, num_lines > 0 ? 4 : 3, prhs);
  long n_out = (long)mxGetNumberOfElements(plhs[0]);
  memcpy(profile_out, plhs[0]->pr, (size_t)n_out * sizeof(double));
  free(plhs[0]->pr);
  free(plhs[0]);
  return n_out;
}
