"""oracle/oracle.py -- TEST INFRASTRUCTURE ONLY.

CPU checker for the hot path.  Three independent restatements, cross-checked in
tests/test_oracle_*.py:

  O1  the reference's voigt.c compiled VERBATIM (oracle/_ref/libvoigt_ref.so, built by
      oracle/Makefile from /root/reference/voigt.c against shim/mex.h + shim/cerf.h)
  O2c plain-C restatement (oracle/dla_oracle.c + cerf_port.c -> _build/liboracle.so)
  O2n numpy restatement below (scipy.special.wofz = the Faddeeva package libcerf 1.x wraps;
      scipy.linalg for chol / triangular solves), following
        voigt.c:278-299, log_mvnpdf_low_rank.m:5-34 (as written),
        process_qsos.m:150-155, :163-165, :184-212, :223-232, generate_ascii_catalog.m:73
  O3  dense brute force  log N(y; mu, M M' + diag d)  by an n x n Cholesky.

Parity status: the reference tree has no tests, goldens or fixtures for this path and
Octave/MATLAB/libcerf are absent, so O1 (voigt.c verbatim, with OUR libcerf stand-in) is
the only piece of the reference that runs here: parity is pinned for voigt.c's own
arithmetic and UNPINNED at the libcerf boundary and for the two .m files.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_C = os.path.join(HERE, "_build", "liboracle.so")
LIB_REF = os.path.join(HERE, "_ref", "libvoigt_ref.so")

_pd = C.POINTER(C.c_double)
_VOIGT_FN = C.CFUNCTYPE(C.c_long, _pd, C.c_long, C.c_double, C.c_double, C.c_int, _pd)

_c = None
_ref = None

# voigt.c literals for the three lines the pipeline uses (voigt.c:22, :33-35, :146, :153-155,
# :189-191, :244-250), for the numpy restatement.
C_CGS = 2.99792458e+10
SIGMA = 9.08537121627923800e+05
TRANSITION_WAVELENGTHS = np.array([1.2156701e-05, 1.0257223e-05, 9.725368e-06])
LEADING_CONSTANTS = np.array([1.34347262962625339e-07, 2.15386482180851912e-08, 7.48525170087141461e-09])
GAMMAS = np.array([6.06075804241938613e+02, 1.54841462408931704e+02, 6.28964942715328164e+01])
INSTRUMENT_PROFILE = np.array([2.17460992138080811e-03, 4.11623059580451742e-02, 2.40309364651846963e-01,
                               4.32707438937454059e-01, 2.40309364651846963e-01, 4.11623059580451742e-02,
                               2.17460992138080811e-03])
LOG_2PI = 1.83787706640934534


def build(force: bool = False):
    """Compile the C restatement and, where /root/reference exists, voigt.c verbatim."""
    if force or not os.path.exists(LIB_C) or (os.path.exists("/root/reference/voigt.c") and not os.path.exists(LIB_REF)):
        env = dict(os.environ)
        env.pop("CC", None)
        subprocess.run(["make", "-C", HERE] + (["-B"] if force else []), check=True, capture_output=True, env=env)


def load_c() -> C.CDLL:
    global _c
    if _c is None:
        build()
        lib = C.CDLL(LIB_C)
        lib.oracle_w_of_z.argtypes = [C.c_double, C.c_double, _pd, _pd]
        lib.voigt.restype = C.c_double
        lib.voigt.argtypes = [C.c_double] * 3
        lib.oracle_voigt.restype = C.c_long
        lib.oracle_voigt.argtypes = [_pd, C.c_long, C.c_double, C.c_double, C.c_int, _pd]
        lib.oracle_log_mvnpdf_low_rank.restype = C.c_double
        lib.oracle_log_mvnpdf_low_rank.argtypes = [_pd, _pd, _pd, _pd, C.c_long, C.c_int]
        lib.oracle_log_mvnpdf_low_rank_ld.restype = C.c_double
        lib.oracle_log_mvnpdf_low_rank_ld.argtypes = [_pd, _pd, _pd, _pd, C.c_long, C.c_int]
        lib.oracle_process_quasar.restype = C.c_long
        lib.oracle_process_quasar.argtypes = [C.c_long, C.c_long, _pd, C.POINTER(C.c_int), _pd, _pd, _pd, _pd, _pd,
                                              C.c_double, C.c_double, C.c_long, _pd, _pd, C.c_int, C.c_int,
                                              C.c_double, C.c_double, C.c_long, C.c_long, C.c_void_p, C.c_int,
                                              _pd, _pd]
        lib.oracle_max_threads.restype = C.c_int
        _c = lib
    return _c


def load_ref():
    """The reference's voigt.c compiled verbatim, or None if it was never built."""
    global _ref
    if _ref is None:
        build()
        if not os.path.exists(LIB_REF):
            return None
        lib = C.CDLL(LIB_REF)
        lib.ref_voigt.restype = C.c_long
        lib.ref_voigt.argtypes = [_pd, C.c_long, C.c_double, C.c_double, C.c_int, _pd]
        _ref = lib
    return _ref


def _p(a):
    return a.ctypes.data_as(_pd)


def w_of_z_c(x, y):
    a, b = C.c_double(), C.c_double()
    load_c().oracle_w_of_z(float(x), float(y), C.byref(a), C.byref(b))
    return complex(a.value, b.value)


def voigt_c(lambdas, z, N, num_lines=3):
    lam = np.ascontiguousarray(lambdas, dtype=np.float64)
    out = np.empty(lam.shape[0] - 6)
    load_c().oracle_voigt(_p(lam), lam.shape[0], float(z), float(N), int(num_lines), _p(out))
    return out


def voigt_ref(lambdas, z, N, num_lines=3):
    lib = load_ref()
    if lib is None:
        raise RuntimeError("oracle/_ref/libvoigt_ref.so not built (needs /root/reference/voigt.c)")
    lam = np.ascontiguousarray(lambdas, dtype=np.float64)
    out = np.empty(lam.shape[0] - 6)
    lib.ref_voigt(_p(lam), lam.shape[0], float(z), float(N), int(num_lines), _p(out))
    return out


def voigt_numpy(lambdas, z, N, num_lines=3):
    """voigt.c:278-299 with scipy.special.wofz standing in for libcerf."""
    from scipy.special import wofz
    lam = np.asarray(lambdas, dtype=np.float64)
    mult = C_CGS / (TRANSITION_WAVELENGTHS[:num_lines] * (1 + z)) / 1e8          # :278-279
    total = np.zeros(lam.shape[0])
    for j in range(num_lines):                                                    # :285-289
        velocity = lam * mult[j] - C_CGS
        zc = (velocity + 1j * GAMMAS[j]) / np.sqrt(2.0) / SIGMA
        total += -LEADING_CONSTANTS[j] * (wofz(zc).real / np.sqrt(2 * np.pi) / SIGMA)
    raw = np.exp(N * total)                                                       # :291
    prof = np.zeros(lam.shape[0] - 6)
    for t in range(7):                                                            # :297-299, tap order
        prof += raw[t:t + prof.shape[0]] * INSTRUMENT_PROFILE[t]
    return prof


def log_mvnpdf_low_rank_numpy(y, mu, M, d):
    """log_mvnpdf_low_rank.m:5-34, statement by statement."""
    from scipy.linalg import cholesky, solve_triangular
    n, k = M.shape
    y = y - mu                                              # :11
    d_inv = 1.0 / d                                         # :13
    D_inv_y = d_inv * y                                     # :14
    D_inv_M = d_inv[:, None] * M                            # :15
    B = M.T @ D_inv_M                                       # :22
    B[np.diag_indices(k)] += 1                              # :23
    L = cholesky(B, lower=False)                            # :24  L'L = B
    Cm = solve_triangular(L, solve_triangular(L, D_inv_M.T, trans="T", lower=False), lower=False)  # :26
    K_inv_y = D_inv_y - D_inv_M @ (Cm @ y)                  # :28
    log_det_K = np.sum(np.log(d)) + 2 * np.sum(np.log(np.diag(L)))  # :30
    return -0.5 * (y @ K_inv_y + log_det_K + n * LOG_2PI)   # :32


def log_mvnpdf_low_rank_c(y, mu, M, d):
    n, k = M.shape
    Mf = np.ascontiguousarray(np.asfortranarray(M).ravel(order="F"))
    y, mu, d = (np.ascontiguousarray(a, dtype=np.float64) for a in (y, mu, d))
    return load_c().oracle_log_mvnpdf_low_rank(_p(y), _p(mu), _p(Mf), _p(d), n, k)


def log_mvnpdf_low_rank_extended(y, mu, M, d):
    """The same statements in x87 extended precision (oracle/lmlr_extended.c): how far the reference's fp64
    evaluation is from the exact value of its own formula (tools/parity_truth.py)."""
    n, k = M.shape
    Mf = np.ascontiguousarray(np.asfortranarray(M).ravel(order="F"))
    y, mu, d = (np.ascontiguousarray(a, dtype=np.float64) for a in (y, mu, d))
    return load_c().oracle_log_mvnpdf_low_rank_ld(_p(y), _p(mu), _p(Mf), _p(d), n, k)


def log_mvnpdf_dense(y, mu, M, d):
    """O3: the definition, by an n x n Cholesky (independent of the Woodbury algebra)."""
    from scipy.linalg import cho_solve, cholesky
    n = y.shape[0]
    K = M @ M.T + np.diag(d)
    Lc = cholesky(K, lower=True)
    r = y - mu
    alpha = cho_solve((Lc, True), r)
    return -0.5 * (r @ alpha + 2 * np.sum(np.log(np.diag(Lc))) + n * LOG_2PI)


def reduce_samples(sample_ll, ll_no_dla, log_prior_dla, log_prior_no_dla):
    """process_qsos.m:202-212, :223-229 and generate_ascii_catalog.m:73 on one row."""
    ll = np.asarray(sample_ll, dtype=np.float64)
    finite = ~np.isnan(ll)
    m = np.max(ll[finite]) if finite.any() else np.nan                 # MATLAB max ignores NaN
    with np.errstate(invalid="ignore", divide="ignore"):
        ll_dla = m + np.log(np.mean(np.exp(ll - m)))                   # :204-209
    map_ind = int(np.flatnonzero(finite & (ll == m))[0]) if finite.any() else -1
    a = log_prior_no_dla + ll_no_dla                                   # :154-155
    b = log_prior_dla + ll_dla                                         # :211-212
    mx = np.nanmax([a, b])
    e = np.exp(np.array([a, b]) - mx)
    return {"log_likelihoods_dla": ll_dla, "log_posteriors_no_dla": a, "log_posteriors_dla": b,
            "model_posteriors": e / e.sum(), "map_ind": map_ind}


def process_quasar_c(prep, offset_samples, nhi_samples, log_prior_dla, log_prior_no_dla, num_lines=3,
                     use_ref_voigt=False, nthreads=0, first_sample=0, num_eval=0):
    """One quasar through the C restatement (optionally with voigt.c verbatim for the profile)."""
    lib = load_c()
    S = offset_samples.shape[0]
    k = prep.M.shape[1]
    Mf = np.ascontiguousarray(np.asfortranarray(prep.M).ravel(order="F"))
    pidx = np.ascontiguousarray(prep.profile_index, dtype=np.int32)
    off = np.ascontiguousarray(offset_samples, dtype=np.float64)
    nhi = np.ascontiguousarray(nhi_samples, dtype=np.float64)
    sample_ll = np.full(S, np.nan)
    scal = np.zeros(6)
    fn = None
    if use_ref_voigt:
        ref = load_ref()
        if ref is None:
            raise RuntimeError("libvoigt_ref.so not built")
        fn = C.cast(ref.ref_voigt, C.c_void_p)
    if nthreads <= 0:
        nthreads = lib.oracle_max_threads()
    mi = lib.oracle_process_quasar(prep.n, prep.n_full, _p(prep.padded_wavelengths),
                                   pidx.ctypes.data_as(C.POINTER(C.c_int)), _p(prep.flux), _p(prep.mu), _p(Mf),
                                   _p(prep.omega2), _p(prep.noise_variance), prep.min_z_dla, prep.max_z_dla, S,
                                   _p(off), _p(nhi), int(num_lines), k, float(log_prior_dla),
                                   float(log_prior_no_dla), int(first_sample), int(num_eval), fn, int(nthreads),
                                   _p(sample_ll), _p(scal))
    return {"sample_log_likelihoods_dla": sample_ll, "log_likelihoods_no_dla": scal[0],
            "log_likelihoods_dla": scal[1], "log_posteriors_no_dla": scal[2], "log_posteriors_dla": scal[3],
            "model_posteriors": scal[4:6].copy(), "map_ind": int(mi)}


def process_quasar_numpy(prep, offset_samples, nhi_samples, log_prior_dla, log_prior_no_dla, num_lines=3):
    """One quasar through the numpy restatement (slow: use on small cases)."""
    S = offset_samples.shape[0]
    ll_no = log_mvnpdf_low_rank_numpy(prep.flux, prep.mu, np.asarray(prep.M), prep.omega2 + prep.noise_variance)
    sample_z = prep.min_z_dla + (prep.max_z_dla - prep.min_z_dla) * offset_samples          # :163-165
    ll = np.empty(S)
    for i in range(S):                                                                        # :184
        absorption = voigt_numpy(prep.padded_wavelengths, sample_z[i], nhi_samples[i], num_lines)  # :186
        absorption = absorption[prep.profile_index]                                           # :189
        dla_mu = prep.mu * absorption                                                         # :191
        dla_M = np.asarray(prep.M) * absorption[:, None]                                      # :192
        dla_omega2 = prep.omega2 * absorption ** 2                                            # :193
        ll[i] = log_mvnpdf_low_rank_numpy(prep.flux, dla_mu, dla_M, dla_omega2 + prep.noise_variance)  # :195-197
    out = reduce_samples(ll, ll_no, log_prior_dla, log_prior_no_dla)
    out["sample_log_likelihoods_dla"] = ll
    out["log_likelihoods_no_dla"] = ll_no
    return out
