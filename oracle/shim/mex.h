/* oracle/shim/mex.h -- TEST INFRASTRUCTURE ONLY (not part of the product).
 *
 * A minimal stand-in for MATLAB/Octave's mex.h so that the reference's
 * /root/reference/voigt.c can be compiled VERBATIM as plain C (no MATLAB, no
 * Octave in this image).  It provides exactly the nine symbols voigt.c uses
 * (voigt.c:253-277, :294, :301-302): mxArray, mwSize, mxREAL, mxGetPr,
 * mxGetScalar, mxGetNumberOfElements, mxCreateDoubleMatrix (zero-initialised,
 * which voigt.c:299's "+=" relies on), mxMalloc, mxFree.
 */
#ifndef ORACLE_SHIM_MEX_H
#define ORACLE_SHIM_MEX_H
#include <stddef.h>
#include <stdlib.h>

typedef size_t mwSize;
typedef enum { mxREAL = 0, mxCOMPLEX = 1 } mxComplexity;

typedef unsigned char mxLogical;
typedef struct mxArray_shim {
  double *pr;
  mwSize m, n;
  struct mxArray_shim **cells;   /* cell arrays (dla_process_mex.c): m * n elements */
  mxLogical *logicals;           /* logical arrays */
} mxArray;

static inline double *mxGetPr(const mxArray *a) { return a->pr; }
static inline double mxGetScalar(const mxArray *a) { return a->pr[0]; }
static inline mwSize mxGetNumberOfElements(const mxArray *a) { return a->m * a->n; }
static inline mxArray *mxCreateDoubleMatrix(mwSize m, mwSize n, mxComplexity c) {
  (void)c;
  mxArray *a = (mxArray *)calloc(1, sizeof(mxArray));
  a->m = m; a->n = n;
  a->pr = (double *)calloc((m * n) > 0 ? (m * n) : 1, sizeof(double));
  return a;
}
static inline void *mxMalloc(size_t n) { return malloc(n); }
static inline void mxFree(void *p) { free(p); }

/* Extra symbols used only by the product's MEX shims (gp_dla_detection_b200/mex/), so that the
 * CPU test-suite can syntax/link-check them without MATLAB or Octave. */
static inline mwSize mxGetM(const mxArray *a) { return a->m; }
static inline mwSize mxGetN(const mxArray *a) { return a->n; }
static inline mxArray *mxCreateDoubleScalar(double v) {
  mxArray *a = mxCreateDoubleMatrix(1, 1, mxREAL);
  a->pr[0] = v;
  return a;
}
static inline int mxIsDouble(const mxArray *a) { return a->pr != 0; }
static inline const mxArray *mxGetCell(const mxArray *a, mwSize i) { return a->cells ? a->cells[i] : 0; }
static inline int mxIsLogical(const mxArray *a) { return a->logicals != 0; }
static inline mxLogical *mxGetLogicals(const mxArray *a) { return a->logicals; }
static inline int mxIsComplex(const mxArray *a) { (void)a; return 0; }
void mexErrMsgIdAndTxt(const char *id, const char *fmt, ...);
int mexAtExit(void (*fn)(void));

void mexFunction(int nlhs, mxArray *plhs[], int nrhs, const mxArray *prhs[]);
#endif
