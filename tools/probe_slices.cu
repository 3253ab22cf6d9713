// tools/probe_slices.cu -- VERDICT r1 item 2: can the Woodbury contraction leave the FP64 pipe?
// Exact-slice int8 tcgen05 contraction (gp_dla_detection_b200/csrc/slices.cuh) on operands shaped like a config-2
// quasar (1217 -> 1248 pixels, k = 20: 210 vech columns + 20 M columns), against
//   * a double-double accumulated reference (the "truth"),
//   * a plain fp64 FMA-chain dot product (what the DMMA path's accumulation is worth),
// for NS = 6, 7, 8 slices; then the throughput of the contraction alone on 16 quasars' worth of tiles.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tools/_bin/probe_slices tools/probe_slices.cu
// Output: one JSON object per line (-> profiles/r02_probe_slices.jsonl).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../gp_dla_detection_b200/csrc/slices.cuh"

using namespace dla::sl;

#define CK(x)                                                                                   \
  do {                                                                                          \
    cudaError_t e_ = (x);                                                                       \
    if (e_ != cudaSuccess) {                                                                    \
      fprintf(stderr, "CUDA error %s at %s:%d: %s\n", #x, __FILE__, __LINE__, cudaGetErrorString(e_)); \
      exit(2);                                                                                  \
    }                                                                                           \
  } while (0)

// ---- references ----------------------------------------------------------------------------------------
__global__ void ref_kernel(const double* __restrict__ X, int64_t ldx, const double* __restrict__ T, int64_t ldt, int rows, int npix,
                           int ncol, double* __restrict__ Gdd, double* __restrict__ Gf64, double* __restrict__ Gabs) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x, r = blockIdx.y;
  if (c >= ncol || r >= rows) return;
  double hi = 0.0, lo = 0.0, f = 0.0, ab = 0.0;
  for (int p = 0; p < npix; ++p) {
    const double x = X[(int64_t)r * ldx + p], t = T[(int64_t)p * ldt + c];
    const double ph = x * t, pl = fma(x, t, -ph);          // TwoProd
    const double s = hi + ph, bb = s - hi;                 // TwoSum
    const double err = (hi - (s - bb)) + (ph - bb);
    hi = s;
    lo += err + pl;
    f = fma(x, t, f);
    ab += fabs(x) * fabs(t);
  }
  Gdd[(int64_t)r * ncol + c] = hi + lo;
  Gf64[(int64_t)r * ncol + c] = f;
  Gabs[(int64_t)r * ncol + c] = ab;
}

// F2I.S64.F64 issue rate (is the conversion K1 would need cheap?)
__global__ void f2i_probe(long long* out, int iters, double x0) {
  double x[8];
  long long acc = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) x[i] = x0 + threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      acc += __double2ll_rn(x[i]);
      x[i] += 1.0;
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

struct Lcg {
  uint64_t s;
  double u() { s = s * 6364136223846793005ull + 1442695040888963407ull; return ((s >> 11) + 0.5) / 9007199254740992.0; }
  double n() { const double a = u(), b = u(); return sqrt(-2 * log(a)) * cos(6.283185307179586 * b); }
};

template <int NS, int ORDER>
static float run_contract(const SliceJob* d_jobs, int njobs, const int* d_tp, int reps) {
  CK(cudaFuncSetAttribute(contract_slices_kernel<NS, ORDER>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Geo<NS>::SMEM));
  int sms;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int r = 0; r < reps; ++r) {
    CK(cudaEventRecord(e0));
    contract_slices_kernel<NS, ORDER><<<sms, THREADS, Geo<NS>::SMEM>>>(d_jobs, njobs, d_tp);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    CK(cudaGetLastError());
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    if (ms < best) best = ms;
  }
  return best;
}

template <int NS>
static void accuracy_and_speed(const std::vector<double>& hW, const std::vector<double>& hU, const std::vector<double>& hT,
                               const std::vector<double>& hM, const std::vector<double>& icw, const std::vector<double>& cw,
                               const std::vector<double>& icu, const std::vector<double>& cu, int rows, int npix, int k,
                               int time_quasars) {
  const int T = k * (k + 1) / 2, ncolW = 256, ncolU = 64;   // padded column counts (4 + 1 column blocks)
  double *dW, *dU, *dT, *dM, *dicw, *dcw, *dicu, *dcu;
  CK(cudaMalloc(&dW, hW.size() * 8));  CK(cudaMemcpy(dW, hW.data(), hW.size() * 8, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&dU, hU.size() * 8));  CK(cudaMemcpy(dU, hU.data(), hU.size() * 8, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&dT, hT.size() * 8));  CK(cudaMemcpy(dT, hT.data(), hT.size() * 8, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&dM, hM.size() * 8));  CK(cudaMemcpy(dM, hM.data(), hM.size() * 8, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&dicw, npix * 8));     CK(cudaMemcpy(dicw, icw.data(), npix * 8, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&dcw, npix * 8));      CK(cudaMemcpy(dcw, cw.data(), npix * 8, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&dicu, npix * 8));     CK(cudaMemcpy(dicu, icu.data(), npix * 8, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&dcu, npix * 8));      CK(cudaMemcpy(dcu, cu.data(), npix * 8, cudaMemcpyHostToDevice));
  const int nks = npix / KS, nsb = rows / SB;
  int8_t *pAW, *pAU, *pBW, *pBU;
  CK(cudaMalloc(&pAW, (size_t)nsb * nks * Geo<NS>::A_STEP));
  CK(cudaMalloc(&pAU, (size_t)nsb * nks * Geo<NS>::A_STEP));
  CK(cudaMalloc(&pBW, (size_t)(ncolW / NB) * nks * Geo<NS>::B_STEP));
  CK(cudaMalloc(&pBU, (size_t)(ncolU / NB) * nks * Geo<NS>::B_STEP));
  double *tauW, *csW, *tauU, *csU, *GW, *GU;
  CK(cudaMalloc(&tauW, ncolW * 8)); CK(cudaMalloc(&csW, ncolW * 8));
  CK(cudaMalloc(&tauU, ncolU * 8)); CK(cudaMalloc(&csU, ncolU * 8));
  CK(cudaMalloc(&GW, (size_t)rows * ncolW * 8));
  CK(cudaMalloc(&GU, (size_t)rows * ncolU * 8));
  CK(cudaMemset(GW, 0xff, (size_t)rows * ncolW * 8));
  CK(cudaMemset(GU, 0xff, (size_t)rows * ncolU * 8));
  // slice the operands
  {
    const int64_t na = (int64_t)rows * (npix / 16);
    slice_rows_kernel<NS><<<(unsigned)((na + 255) / 256), 256>>>(dW, npix, dicw, rows, npix, pAW);
    slice_rows_kernel<NS><<<(unsigned)((na + 255) / 256), 256>>>(dU, npix, dicu, rows, npix, pAU);
    column_scale_kernel<<<ncolW, 256>>>(dT, ncolW, dcw, npix, ncolW, tauW, csW);
    column_scale_kernel<<<ncolU, 256>>>(dM, ncolU, dcu, npix, ncolU, tauU, csU);
    const int64_t nbw = (int64_t)ncolW * (npix / 16), nbu = (int64_t)ncolU * (npix / 16);
    slice_table_kernel<NS><<<(unsigned)((nbw + 255) / 256), 256>>>(dT, ncolW, dcw, tauW, npix, ncolW, pBW);
    slice_table_kernel<NS><<<(unsigned)((nbu + 255) / 256), 256>>>(dM, ncolU, dcu, tauU, npix, ncolU, pBU);
    CK(cudaDeviceSynchronize());
  }
  SliceJob hj[2] = {{pAW, pBW, csW, GW, (int64_t)nks * Geo<NS>::A_STEP, (int64_t)nks * Geo<NS>::B_STEP, nsb, ncolW / NB, nks, ncolW},
                    {pAU, pBU, csU, GU, (int64_t)nks * Geo<NS>::A_STEP, (int64_t)nks * Geo<NS>::B_STEP, nsb, ncolU / NB, nks, ncolU}};
  int htp[3] = {0, nsb * (ncolW / NB), nsb * (ncolW / NB) + nsb * (ncolU / NB)};
  SliceJob* dj;
  int* dtp;
  CK(cudaMalloc(&dj, sizeof hj)); CK(cudaMemcpy(dj, hj, sizeof hj, cudaMemcpyHostToDevice));
  CK(cudaMalloc(&dtp, sizeof htp)); CK(cudaMemcpy(dtp, htp, sizeof htp, cudaMemcpyHostToDevice));
  for (int order = 0; order < 2; ++order) {
    CK(cudaMemset(GW, 0xff, (size_t)rows * ncolW * 8));
    CK(cudaMemset(GU, 0xff, (size_t)rows * ncolU * 8));
    if (order == 0) run_contract<NS, 0>(dj, 2, dtp, 1); else run_contract<NS, 1>(dj, 2, dtp, 1);
    CK(cudaDeviceSynchronize());
    // references
    double *rdd, *rf, *rab;
    auto check = [&](const char* name, const double* dX, const double* dTab, int ldt, int ncol_used, const double* dG, int ldg) {
      CK(cudaMalloc(&rdd, (size_t)rows * ncol_used * 8)); CK(cudaMalloc(&rf, (size_t)rows * ncol_used * 8));
      CK(cudaMalloc(&rab, (size_t)rows * ncol_used * 8));
      ref_kernel<<<dim3((ncol_used + 63) / 64, rows), 64>>>(dX, npix, dTab, ldt, rows, npix, ncol_used, rdd, rf, rab);
      CK(cudaDeviceSynchronize());
      std::vector<double> hdd((size_t)rows * ncol_used), hf(hdd.size()), hab(hdd.size()), hg((size_t)rows * ldg);
      CK(cudaMemcpy(hdd.data(), rdd, hdd.size() * 8, cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(hf.data(), rf, hf.size() * 8, cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(hab.data(), rab, hab.size() * 8, cudaMemcpyDeviceToHost));
      CK(cudaMemcpy(hg.data(), dG, hg.size() * 8, cudaMemcpyDeviceToHost));
      double e_sl = 0, e_f = 0, e_sl_rel = 0, e_f_rel = 0, rms_sl = 0, rms_f = 0;
      long bad = 0;
      for (int r = 0; r < rows; ++r)
        for (int c = 0; c < ncol_used; ++c) {
          const double t = hdd[(size_t)r * ncol_used + c], ab = hab[(size_t)r * ncol_used + c];
          const double g = hg[(size_t)r * ldg + c], f = hf[(size_t)r * ncol_used + c];
          if (!(g == g)) { ++bad; continue; }
          const double a = fabs(g - t) / ab, b = fabs(f - t) / ab;
          e_sl = fmax(e_sl, a); e_f = fmax(e_f, b);
          rms_sl += a * a; rms_f += b * b;
          if (fabs(t) > 0) { e_sl_rel = fmax(e_sl_rel, fabs(g - t) / fabs(t)); e_f_rel = fmax(e_f_rel, fabs(f - t) / fabs(t)); }
        }
      const double cnt = (double)rows * ncol_used;
      printf("{\"probe\": \"accuracy\", \"product\": \"%s\", \"slices\": %d, \"order\": %d, \"rows\": %d, \"pixels\": %d, \"columns\": %d, "
             "\"nan_entries\": %ld, \"max_err_over_sum_abs\": {\"slices\": %.3e, \"fp64_fma_chain\": %.3e}, "
             "\"rms_err_over_sum_abs\": {\"slices\": %.3e, \"fp64_fma_chain\": %.3e}, "
             "\"max_err_over_abs_value\": {\"slices\": %.3e, \"fp64_fma_chain\": %.3e}, \"log2_max_slices\": %.1f}\n",
             name, NS, order, rows, npix, ncol_used, bad, e_sl, e_f, sqrt(rms_sl / cnt), sqrt(rms_f / cnt), e_sl_rel, e_f_rel,
             log2(e_sl > 0 ? e_sl : 1e-300));
      fflush(stdout);
      CK(cudaFree(rdd)); CK(cudaFree(rf)); CK(cudaFree(rab));
    };
    check("W.Phi", dW, dT, ncolW, T, GW, ncolW);
    check("U.M", dU, dM, ncolU, k, GU, ncolU);
  }
  // ---- throughput: time_quasars quasars x 10048 samples (79 sample blocks; the planes are reused cyclically so the
  //      A traffic is as large as the real thing but the allocation stays small) ----
  if (time_quasars > 0) {
    const int S_pad = 79 * SB;
    const int nsbq = S_pad / SB;
    int8_t* bigA;
    const size_t per_q = (size_t)nsbq * nks * Geo<NS>::A_STEP;
    CK(cudaMalloc(&bigA, per_q * time_quasars));
    for (size_t off = 0; off < per_q * time_quasars; off += (size_t)nsb * nks * Geo<NS>::A_STEP) {
      const size_t len = std::min((size_t)nsb * nks * Geo<NS>::A_STEP, per_q * time_quasars - off);
      CK(cudaMemcpy(bigA + off, pAW, len, cudaMemcpyDeviceToDevice));
    }
    double* bigG;
    CK(cudaMalloc(&bigG, (size_t)time_quasars * S_pad * ncolW * 8));
    std::vector<SliceJob> jobs;
    std::vector<int> tp{0};
    for (int q = 0; q < time_quasars; ++q) {
      jobs.push_back({bigA + per_q * q, pBW, csW, bigG + (size_t)q * S_pad * ncolW, (int64_t)nks * Geo<NS>::A_STEP,
                      (int64_t)nks * Geo<NS>::B_STEP, nsbq, ncolW / NB, nks, ncolW});
      tp.push_back(tp.back() + nsbq * (ncolW / NB));
    }
    SliceJob* dj2; int* dtp2;
    CK(cudaMalloc(&dj2, jobs.size() * sizeof(SliceJob))); CK(cudaMemcpy(dj2, jobs.data(), jobs.size() * sizeof(SliceJob), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&dtp2, tp.size() * 4)); CK(cudaMemcpy(dtp2, tp.data(), tp.size() * 4, cudaMemcpyHostToDevice));
    for (int order = 0; order < 2; ++order) {
      const float ms = order == 0 ? run_contract<NS, 0>(dj2, time_quasars, dtp2, 4) : run_contract<NS, 1>(dj2, time_quasars, dtp2, 4);
      const double mmas = (double)tp.back() * nks * (NS * (NS + 1) / 2);
      const double int_ops = mmas * 2.0 * SB * NB * KS;
      const double f64_equiv = (double)time_quasars * S_pad * 2.0 * npix * ncolW;       // padded columns
      const double f64_useful = (double)time_quasars * 10001 * 2.0 * 1217 * 230;          // what K2 counts: n k (k+3) per sample
      printf("{\"probe\": \"throughput\", \"slices\": %d, \"order\": %d, \"quasars\": %d, \"tiles\": %d, \"ms\": %.3f, "
             "\"int8_tops\": %.1f, \"fp64_equiv_tflops_padded\": %.1f, \"fp64_equiv_tflops_useful\": %.1f, "
             "\"ms_per_64_quasars\": %.2f, \"a_plane_gbs\": %.0f}\n",
             NS, order, time_quasars, tp.back(), ms, int_ops / ms / 1e9, f64_equiv / ms / 1e9, f64_useful / ms / 1e9,
             ms * 64.0 / time_quasars, (double)per_q * time_quasars / ms / 1e6);
      fflush(stdout);
    }
    CK(cudaFree(bigA)); CK(cudaFree(bigG)); CK(cudaFree(dj2)); CK(cudaFree(dtp2));
  }
  for (void* p : {(void*)dW, (void*)dU, (void*)dT, (void*)dM, (void*)dicw, (void*)dcw, (void*)dicu, (void*)dcu, (void*)pAW, (void*)pAU,
                  (void*)pBW, (void*)pBU, (void*)tauW, (void*)csW, (void*)tauU, (void*)csU, (void*)GW, (void*)GU, (void*)dj, (void*)dtp})
    CK(cudaFree(p));
}

int main(int argc, char** argv) {
  const int time_quasars = argc > 1 ? atoi(argv[1]) : 16;
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz\": %d}\n", prop.name, prop.multiProcessorCount, prop.clockRate);
  const int rows = 256, n = 1217, npix = 1248, k = 20, T = k * (k + 1) / 2;
  // a config-2-like quasar: noise variances lognormal, omega2, a damped trough per sample row, M as synthetic.make_model
  Lcg g{20260002};
  std::vector<double> nu(npix, 1.0), om2(npix, 0.0), mu(npix, 0.0), y(npix, 0.0);
  std::vector<double> M((size_t)npix * k, 0.0);
  for (int p = 0; p < n; ++p) {
    nu[p] = fmin(fmax(exp(log(0.05) + 0.7 * g.n()), 1e-3), 4.0);
    const double lo = 0.12 + 0.08 * pow(sin(3.14159265358979 * p * 0.25 / 90), 2);
    om2[p] = lo * lo;
    mu[p] = 1 + 0.8 * exp(-pow((911.75 + 0.25 * p - 1215.67) / 12, 2)) + 0.1 * (911.75 + 0.25 * p - 1060) / 150;
  }
  std::vector<double> phase(k);
  for (int j = 0; j < k; ++j) phase[j] = 6.283185307179586 * g.u();
  for (int p = 0; p < n; ++p)
    for (int j = 0; j < k; ++j) M[(size_t)p * k + j] = 0.35 * pow(j + 1.0, -0.7) * cos(3.14159265358979 * (j + 0.5) * p / 1216.0 + phase[j]);
  std::vector<double> zeta(k);
  for (int j = 0; j < k; ++j) zeta[j] = g.n();
  for (int p = 0; p < n; ++p) {
    double s = mu[p];
    for (int j = 0; j < k; ++j) s += M[(size_t)p * k + j] * zeta[j];
    y[p] = s + sqrt(om2[p] + nu[p]) * g.n();
  }
  // tables: T (vech of M_p M_p^T, 256 padded columns), Mtab (64 padded columns)
  std::vector<double> hT((size_t)npix * 256, 0.0), hM((size_t)npix * 64, 0.0);
  for (int p = 0; p < n; ++p) {
    int c = 0;
    for (int i = 0; i < k; ++i)
      for (int l = 0; l <= i; ++l) hT[(size_t)p * 256 + c++] = M[(size_t)p * k + i] * M[(size_t)p * k + l];
    for (int j = 0; j < k; ++j) hM[(size_t)p * 64 + j] = M[(size_t)p * k + j];
  }
  (void)T;
  // per-pixel powers of two: cw >= 1/(om2 + nu) >= any w; cu >= (|y| + |mu|) max_a a/(om2 a^2 + nu) >= any |u|
  std::vector<double> cw(npix, 1.0), icw(npix, 1.0), cu(npix, 1.0), icu(npix, 1.0);
  for (int p = 0; p < n; ++p) {
    int e;
    frexp(1.0 / (om2[p] + nu[p]), &e);
    cw[p] = ldexp(1.0, e); icw[p] = ldexp(1.0, -e);
    const double amax = nu[p] < om2[p] ? 0.5 / sqrt(om2[p] * nu[p]) : 1.0 / (om2[p] + nu[p]);
    frexp((fabs(y[p]) + fabs(mu[p])) * amax, &e);
    cu[p] = ldexp(1.0, e); icu[p] = ldexp(1.0, -e);
  }
  // sample rows: row 0 = no DLA (a == 1); others a damped Lorentzian-like trough of random centre / width
  std::vector<double> hW((size_t)rows * npix, 0.0), hU((size_t)rows * npix, 0.0);
  for (int r = 0; r < rows; ++r) {
    const double c0 = g.u() * n, wd = 2.0 + 60.0 * g.u();
    for (int p = 0; p < n; ++p) {
      double a = 1.0;
      if (r > 0) {
        const double x = (p - c0) / wd;
        a = exp(-1.0 / (x * x + 1e-3));
      }
      const double d = om2[p] * a * a + nu[p], rr = y[p] - mu[p] * a;
      hW[(size_t)r * npix + p] = a * a / d;
      hU[(size_t)r * npix + p] = a * rr / d;
    }
  }
  // F2I.S64.F64 issue rate
  {
    long long* out;
    CK(cudaMalloc(&out, (size_t)148 * 8 * 256 * 8));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    f2i_probe<<<148 * 8, 256>>>(out, 100, 1.5);
    CK(cudaEventRecord(e0));
    f2i_probe<<<148 * 8, 256>>>(out, 20000, 1.5);
    CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    const double conv = 148.0 * 8 * 256 * 20000 * 8;
    printf("{\"probe\": \"f2i_s64_f64\", \"ms\": %.3f, \"gconv_per_s\": %.1f, \"per_clk_per_sm\": %.2f}\n", ms, conv / ms / 1e6,
           conv / ms / 1e3 / 148 / (prop.clockRate));
    CK(cudaFree(out));
  }
  accuracy_and_speed<8>(hW, hU, hT, hM, icw, cw, icu, cu, rows, npix, k, time_quasars);
  accuracy_and_speed<7>(hW, hU, hT, hM, icw, cw, icu, cu, rows, npix, k, time_quasars);
  accuracy_and_speed<6>(hW, hU, hT, hM, icw, cw, icu, cu, rows, npix, k, time_quasars);
  return 0;
}
