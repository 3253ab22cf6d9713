/* tests/mex_harness.c -- TEST INFRASTRUCTURE.  Calls the mexFunction of gp_dla_detection_b200/mex/dla_process_mex.c the
 * way MATLAB would, with mxArrays built on the stand-in mex.h (oracle/shim): cell arrays of per-quasar vectors in, the
 * variables of process_qsos.m:235-243 out.  Lets the GPU test-suite EXECUTE the gateway without MATLAB / Octave. */
#include <setjmp.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>
#include "mex.h"

static jmp_buf g_jmp;
static char g_msg[512];
void mexErrMsgIdAndTxt(const char *id, const char *fmt, ...) {
  va_list ap;
  (void)id;
  va_start(ap, fmt);
  vsnprintf(g_msg, sizeof g_msg, fmt, ap);
  va_end(ap);
  longjmp(g_jmp, 1);
}
static void (*g_atexit)(void) = 0;
int mexAtExit(void (*fn)(void)) { g_atexit = fn; return 0; }
const char *harness_last_error(void) { return g_msg; }
void harness_unload(void) { if (g_atexit) g_atexit(); g_atexit = 0; }

static mxArray vec(const double *p, mwSize m, mwSize n) {
  mxArray a;
  memset(&a, 0, sizeof a);
  a.pr = (double *)p; a.m = m; a.n = n;
  return a;
}

int harness_process(long Q, const long *spec_ptr, const double *wl, const double *fl, const double *nv,
                    const unsigned char *mask, int mask_as_logical, const double *z, const double *lpd, const double *lpn,
                    long n_rest, int k, const double *rest, const double *mu, const double *M, const double *lo,
                    double c0, double tau0, double beta, long S, const double *off, const double *nhi,
                    const double *params10, int nlhs, double **outs /* 9 pointers, sizes as the gateway's outputs */) {
  mxArray *cw = calloc(Q > 0 ? Q : 1, sizeof(mxArray)), *cf = calloc(Q > 0 ? Q : 1, sizeof(mxArray));
  mxArray *cv = calloc(Q > 0 ? Q : 1, sizeof(mxArray)), *cm = calloc(Q > 0 ? Q : 1, sizeof(mxArray));
  mxArray **pw = calloc(Q > 0 ? Q : 1, sizeof(mxArray *)), **pf = calloc(Q > 0 ? Q : 1, sizeof(mxArray *));
  mxArray **pv = calloc(Q > 0 ? Q : 1, sizeof(mxArray *)), **pm = calloc(Q > 0 ? Q : 1, sizeof(mxArray *));
  double *maskd = 0;
  long q, total = spec_ptr[Q], i;
  mxArray a[17];
  const mxArray *prhs[17];
  mxArray *plhs[9] = {0};
  int rc = 0, o;
  if (!mask_as_logical) {
    maskd = malloc((size_t)(total > 0 ? total : 1) * sizeof(double));
    for (i = 0; i < total; i++) maskd[i] = mask[i];
  }
  for (q = 0; q < Q; q++) {
    const long s = spec_ptr[q], n = spec_ptr[q + 1] - s;
    cw[q] = vec(wl + s, (mwSize)n, 1); cf[q] = vec(fl + s, (mwSize)n, 1); cv[q] = vec(nv + s, (mwSize)n, 1);
    memset(&cm[q], 0, sizeof(mxArray));
    cm[q].m = (mwSize)n; cm[q].n = 1;
    if (mask_as_logical) cm[q].logicals = (mxLogical *)mask + s; else cm[q].pr = maskd + s;
    pw[q] = &cw[q]; pf[q] = &cf[q]; pv[q] = &cv[q]; pm[q] = &cm[q];
  }
  memset(a, 0, sizeof a);
  a[0].cells = pw; a[1].cells = pf; a[2].cells = pv; a[3].cells = pm;
  for (o = 0; o < 4; o++) { a[o].m = (mwSize)Q; a[o].n = 1; }
  a[4] = vec(z, (mwSize)Q, 1); a[5] = vec(lpd, (mwSize)Q, 1); a[6] = vec(lpn, (mwSize)Q, 1);
  a[7] = vec(rest, (mwSize)n_rest, 1); a[8] = vec(mu, (mwSize)n_rest, 1); a[9] = vec(M, (mwSize)n_rest, (mwSize)k);
  a[10] = vec(lo, (mwSize)n_rest, 1);
  a[11] = vec(&c0, 1, 1); a[12] = vec(&tau0, 1, 1); a[13] = vec(&beta, 1, 1);
  a[14] = vec(off, 1, (mwSize)S); a[15] = vec(nhi, 1, (mwSize)S); a[16] = vec(params10, 1, 10);
  for (o = 0; o < 17; o++) prhs[o] = &a[o];
  if (setjmp(g_jmp) == 0) {
    mexFunction(nlhs, plhs, 17, prhs);
    for (o = 0; o < 9 && o < (nlhs > 0 ? nlhs : 1); o++)
      if (plhs[o] && outs[o]) memcpy(outs[o], plhs[o]->pr, plhs[o]->m * plhs[o]->n * sizeof(double));
  } else {
    rc = -1;
  }
  for (o = 0; o < 9; o++)
    if (plhs[o]) { free(plhs[o]->pr); free(plhs[o]); }
  free(cw); free(cf); free(cv); free(cm); free(pw); free(pf); free(pv); free(pm); free(maskd);
  return rc;
}
