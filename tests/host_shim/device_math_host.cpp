// Host-side compile of the product's DLA_HD device math (faddeeva.cuh) so that the CPU test
// suite can check it against the oracle without a GPU.  Test infrastructure only: nothing in
// the product calls this.
#include "../../gp_dla_detection_b200/csrc/faddeeva.cuh"
#include "../../gp_dla_detection_b200/csrc/faddeeva_table.cuh"

extern "C" void host_re_w_many(const double* x, double y, long n, double* out) {
  for (long i = 0; i < n; ++i) out[i] = dla_re_w(x[i], y, dla_wgrid);
}
