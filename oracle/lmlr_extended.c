/* oracle/lmlr_extended.c -- TEST INFRASTRUCTURE.  log_mvnpdf_low_rank.m:5-34 evaluated in x87 extended precision
 * (long double: 64-bit mantissa, 11 bits more than the reference's own arithmetic), same statements in the same
 * order as oracle_log_mvnpdf_low_rank_w (dla_oracle.c).  Its purpose is to say how far the REFERENCE'S fp64
 * evaluation is from the exact value of its own formula at a given input, which bounds the agreement any two
 * correct fp64 implementations can have (tools/parity_truth.py): the Woodbury form subtracts two quadratic forms of
 * size 1e3..1e5 (:28, :32), so the result carries an absolute rounding noise of order 1e-10 whatever its own size. */
#include <math.h>
#include <stdlib.h>

typedef long double R;

double oracle_log_mvnpdf_low_rank_ld(const double *y_in, const double *mu, const double *M, const double *d, long n,
                                     int k) {
  const R log_2pi = 1.8378770664093454835606594728112353L; /* :7 */
  R *w = (R *)malloc(sizeof(R) * (size_t)(2 * n * k + 2 * k * k + 4 * n + 2 * k + 16));
  R *y = w, *D_inv_y = y + n, *D_inv_M = D_inv_y + n, *C = D_inv_M + n * k, *B = C + n * k, *Cy = B + k * k,
    *K_inv_y = Cy + k;
  long p;
  int i, j, m;
  R log_det_K = 0, quad = 0, s;
  double out;
  if (!w) return NAN;
  for (p = 0; p < n; p++) {
    const R d_inv = (R)1 / (R)d[p]; /* :13 */
    y[p] = (R)y_in[p] - (R)mu[p];   /* :11 */
    D_inv_y[p] = d_inv * y[p];      /* :14 */
    for (j = 0; j < k; j++) D_inv_M[j * n + p] = d_inv * (R)M[j * n + p]; /* :15 */
  }
  for (i = 0; i < k; i++) /* :22 */
    for (j = 0; j < k; j++) {
      s = 0;
      for (p = 0; p < n; p++) s += (R)M[i * n + p] * D_inv_M[j * n + p];
      B[j * k + i] = s;
    }
  for (i = 0; i < k; i++) B[i * k + i] += 1; /* :23 */
  for (j = 0; j < k; j++) {                  /* :24 chol, upper triangle in B */
    s = B[j * k + j];
    for (m = 0; m < j; m++) s -= B[j * k + m] * B[j * k + m];
    if (!(s > 0)) { free(w); return NAN; }
    s = sqrtl(s);
    B[j * k + j] = s;
    for (i = j + 1; i < k; i++) {
      R t = B[i * k + j];
      for (m = 0; m < j; m++) t -= B[j * k + m] * B[i * k + m];
      B[i * k + j] = t / s;
    }
  }
#define L_(r, c_) B[(c_) * k + (r)]
  for (p = 0; p < n; p++) { /* :26 */
    R *col = C + p * k;
    for (i = 0; i < k; i++) {
      R t = D_inv_M[i * n + p];
      for (m = 0; m < i; m++) t -= L_(m, i) * col[m];
      col[i] = t / L_(i, i);
    }
    for (i = k - 1; i >= 0; i--) {
      R t = col[i];
      for (m = i + 1; m < k; m++) t -= L_(i, m) * col[m];
      col[i] = t / L_(i, i);
    }
  }
  for (i = 0; i < k; i++) Cy[i] = 0; /* :28 */
  for (p = 0; p < n; p++)
    for (i = 0; i < k; i++) Cy[i] += C[p * k + i] * y[p];
  for (p = 0; p < n; p++) {
    s = 0;
    for (j = 0; j < k; j++) s += D_inv_M[j * n + p] * Cy[j];
    K_inv_y[p] = D_inv_y[p] - s;
  }
  for (p = 0; p < n; p++) log_det_K += logl((R)d[p]); /* :30 */
  s = 0;
  for (i = 0; i < k; i++) s += logl(L_(i, i));
  log_det_K += 2 * s;
#undef L_
  for (p = 0; p < n; p++) quad += y[p] * K_inv_y[p];
  out = (double)(-0.5L * (quad + log_det_K + (R)n * log_2pi)); /* :32 */
  free(w);
  return out;
}
