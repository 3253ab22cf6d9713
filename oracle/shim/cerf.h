/* oracle/shim/cerf.h -- TEST INFRASTRUCTURE ONLY.
 * Stand-in for libcerf's <cerf.h> (not installed, not vendored, version not
 * pinned by the reference: only "-lcerf", README.md:218).  voigt.c uses one
 * symbol from it (voigt.c:5 include, voigt.c:288 call):
 *     double voigt(double x, double sigma, double gamma);
 * implemented in oracle/cerf_port.c from the published algorithms.
 */
#ifndef ORACLE_SHIM_CERF_H
#define ORACLE_SHIM_CERF_H
double voigt(double x, double sigma, double gamma);
#endif
