"""CPU suite: the oracle pinned against the known-answer tests of SURVEY.md 8(c) (K1-K12) and
against the committed golden vectors (tests/golden/golden_v1.npz, made by make_golden.py from the
reference's voigt.c compiled verbatim + scipy wofz)."""
import os

import numpy as np
import pytest

from tests.golden.make_golden import checksum
from tests.helpers import prepared_config

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_v1.npz"))
SIGMA = 9.08537121627923800e+05


def test_K1_constants_reproduce_their_comment_formulas(oracle):
    """voigt.c:141-149, :186, :223-240: sigma, leading constants, gammas, instrument profile."""
    k, T, m_p, m_e, c = 1.38064852e-16, 1e4, 1.672621898e-24, 9.10938356e-28, 2.99792458e10
    e = 4.803204672997660e-10
    b = np.sqrt(2 * k * T / m_p)
    assert abs(b / np.sqrt(2) - SIGMA) / SIGMA < 4e-16
    f = np.array([0.416400, 0.079120, 0.029000])                      # oscillator strengths (voigt.c:66-69)
    lc = np.pi * e * e * f * oracle.TRANSITION_WAVELENGTHS / (m_e * c)
    assert np.all(np.abs(lc / oracle.LEADING_CONSTANTS - 1) < 1e-14)
    R, ps = 2000.0, 1e-4
    pixel_sigma = 1 / (R * 2 * np.sqrt(2 * np.log(2)) * (10 ** ps - 1))
    ip = np.exp(-0.5 * np.arange(-3, 4) ** 2 / pixel_sigma ** 2)
    ip /= ip.sum()
    assert np.all(np.abs(ip / oracle.INSTRUMENT_PROFILE - 1) < 2e-15)   # taps 0/6 differ in the 16th digit: literal wins


def test_K2_zero_column_density_gives_exactly_one(oracle):
    lam = GOLD["voigt_lambdas"]
    for fn in (oracle.voigt_c, oracle.voigt_numpy) + ((oracle.voigt_ref,) if oracle.load_ref() else ()):
        assert np.all(fn(lam, 3.0, 0.0, 3) == 1.0)


def test_K3_libcerf_voigt_known_values(oracle):
    lib = oracle.load_c()
    g = oracle.GAMMAS[0]
    want = {0: 4.38870376e-07, 1e5: 4.36221415e-07, 1e6: 2.39565188e-07, 5e6: 8.76784644e-12,
            1e8: 1.92967713e-14, 3e9: 2.14355526e-17}
    for x, v in want.items():
        assert abs(lib.voigt(float(x), SIGMA, g) / v - 1) < 2e-8


def test_K4_length_and_locality(oracle):
    lam = GOLD["voigt_lambdas"]
    z, N = float(GOLD["voigt_z"][2]), float(GOLD["voigt_N"][2])
    full = oracle.voigt_c(lam, z, N, 3)
    assert full.shape[0] == lam.shape[0] - 6
    sub = oracle.voigt_c(lam[100:140], z, N, 3)          # output i depends on inputs i..i+6 only
    assert np.array_equal(sub, full[100:134])


def test_faddeeva_port_matches_golden_wofz(oracle):
    for yi, y in enumerate(GOLD["fadd_y"]):
        got = np.array([oracle.w_of_z_c(x, y).real for x in GOLD["fadd_x"]])
        assert np.max(np.abs(got / GOLD["fadd_re_w"][yi] - 1)) < 2e-15


def test_faddeeva_port_matches_mpmath(oracle):
    import mpmath as mp
    mp.mp.dps = 50
    for x, y in [(0.0, 4.7e-4), (0.3, 1.2e-4), (2.5, 4.9e-5), (5.9, 4.7e-4), (6.1, 4.7e-4), (7.99, 1e-4),
                 (8.01, 1e-4), (9.9, 4.9e-5), (30.0, 4.7e-4), (4001.0, 4.7e-4), (1.0, 0.5), (0.2, 7.5), (12.0, 1e-12)]:
        z = mp.mpf(x) + 1j * mp.mpf(y)
        t = mp.exp(-z * z) * mp.erfc(-1j * z)
        w = oracle.w_of_z_c(x, y)
        assert abs(w.real - float(mp.re(t))) <= 1e-13 * float(mp.re(t)), (x, y)   # the Faddeeva package claims 1e-13
        assert abs(w.imag - float(mp.im(t))) <= 1e-13 * abs(float(mp.im(t))) + 1e-300, (x, y)


def test_voigt_restatements_match_reference_voigt_c_golden(oracle):
    """C restatement == voigt.c verbatim bit for bit; numpy (scipy wofz) within 1e-15."""
    lam = GOLD["voigt_lambdas"]
    for i in range(6):
        z, N = float(GOLD["voigt_z"][i]), float(GOLD["voigt_N"][i])
        assert np.array_equal(oracle.voigt_c(lam, z, N, 3), GOLD["voigt_profiles"][i])
        assert np.max(np.abs(oracle.voigt_numpy(lam, z, N, 3) - GOLD["voigt_profiles"][i])) < 1e-15
        if oracle.load_ref() is not None:
            assert np.array_equal(oracle.voigt_ref(lam, z, N, 3), GOLD["voigt_profiles"][i])
    assert np.array_equal(oracle.voigt_c(lam, float(GOLD["voigt_z"][3]), 1e21, 31), GOLD["voigt_profile_31"])


def _rand_case(n, k, seed):
    rng = np.random.default_rng(seed)
    return (rng.standard_normal(n), 0.1 * rng.standard_normal(n), 0.5 * rng.standard_normal((n, k)),
            0.05 * np.exp(0.7 * rng.standard_normal(n)))


def test_K5_closed_forms(oracle):
    y, mu, M, d = _rand_case(120, 3, 1)
    r = y - mu
    assert oracle.log_mvnpdf_low_rank_c(y, mu, 0 * M, d) == pytest.approx(
        -0.5 * (np.sum(r * r / d) + np.sum(np.log(d)) + 120 * oracle.LOG_2PI), rel=1e-14)
    m = M[:, :1]
    s = 1 + np.sum(m[:, 0] ** 2 / d)                                   # Sherman-Morrison, k = 1
    quad = np.sum(r * r / d) - np.sum(m[:, 0] * r / d) ** 2 / s
    ref = -0.5 * (quad + np.sum(np.log(d)) + np.log(s) + 120 * oracle.LOG_2PI)
    assert oracle.log_mvnpdf_low_rank_c(y, mu, m, d) == pytest.approx(ref, rel=1e-13)
    assert oracle.log_mvnpdf_low_rank_numpy(y, mu, m, d) == pytest.approx(ref, rel=1e-13)


@pytest.mark.parametrize("n,k", [(32, 1), (97, 5), (256, 20), (200, 40)])
def test_K6_woodbury_equals_dense_definition(oracle, n, k):
    y, mu, M, d = _rand_case(n, k, n + k)
    dense = oracle.log_mvnpdf_dense(y, mu, M, d)
    assert oracle.log_mvnpdf_low_rank_c(y, mu, M, d) == pytest.approx(dense, rel=1e-10)
    assert oracle.log_mvnpdf_low_rank_numpy(y, mu, M, d) == pytest.approx(dense, rel=1e-10)


def test_K11_recast_form_equals_as_written(oracle):
    """q0 - t^T t, ld + 2 sum log L_jj (what the kernels compute) vs the .m file as written,
    including a low-noise stress case."""
    from scipy.linalg import cholesky, solve_triangular
    for seed, noise in [(3, 0.05), (4, 1e-4)]:
        y, mu, M, d = _rand_case(600, 20, seed)
        d = d * noise / 0.05
        r = y - mu
        B = np.eye(20) + M.T @ (M / d[:, None])
        L = cholesky(B, lower=True)
        t = solve_triangular(L, M.T @ (r / d), lower=True)
        recast = -0.5 * (np.sum(r * r / d) - t @ t + np.sum(np.log(d)) + 2 * np.sum(np.log(np.diag(L))) + 600 * oracle.LOG_2PI)
        assert recast == pytest.approx(oracle.log_mvnpdf_low_rank_c(y, mu, M, d), rel=1e-11)


def test_K8_K9_K10_reductions(oracle):
    S = 50
    out = oracle.reduce_samples(np.full(S, -7.25), -8.0, np.log(0.1), np.log(0.9))
    assert out["log_likelihoods_dla"] == pytest.approx(-7.25, abs=1e-14) and out["map_ind"] == 0
    ll = np.full(S, -np.inf); ll[17] = -3.0
    out = oracle.reduce_samples(ll, -8.0, np.log(0.1), np.log(0.9))
    assert out["log_likelihoods_dla"] == pytest.approx(-3.0 - np.log(S), abs=1e-14) and out["map_ind"] == 17
    assert out["model_posteriors"].sum() == pytest.approx(1.0, abs=1e-15)
    assert (out["model_posteriors"][1] > 0.5) == (out["log_posteriors_dla"] > out["log_posteriors_no_dla"])
    rng = np.random.default_rng(5)
    ll = -100 + 30 * rng.standard_normal(S)
    ll[3] = np.nan                                              # NaN: ignored by max, poisons the mean
    assert np.isnan(oracle.reduce_samples(ll, -8.0, -1.0, -1.0)["log_likelihoods_dla"])
    ll[3] = -50.0
    out = oracle.reduce_samples(ll, -8.0, -1.0, -1.0)
    assert np.sum(np.exp(ll - (out["log_likelihoods_dla"] + np.log(S)))) == pytest.approx(1.0, rel=1e-12)  # K10


def test_golden_tiny_and_config1_subset(oracle):
    """C restatement vs the committed goldens (numpy restatement for tiny; C + voigt.c verbatim for config 1)."""
    for name, Q, S, pre, tol in (("tiny", 4, 64, "tiny_", 1e-11), ("config1", 2, 96, "c1_", 1e-13)):
        params, model, samples, sp, prepared = prepared_config(name, Q, S)
        assert checksum(prepared) == str(GOLD[pre + "checksum"]), "synthetic generator drifted: regenerate goldens"
        for i, p in enumerate(prepared):
            r = oracle.process_quasar_c(p, samples.offset_samples, samples.nhi_samples, sp.log_priors_dla[i],
                                        sp.log_priors_no_dla[i], use_ref_voigt=False)
            g = GOLD[pre + "sample_log_likelihoods_dla"][i]
            assert np.max(np.abs(r["sample_log_likelihoods_dla"] - g) / np.abs(g)) < tol
            assert r["map_ind"] == int(GOLD[pre + "map_ind"][i])
            for key in ("log_likelihoods_no_dla", "log_likelihoods_dla", "log_posteriors_dla"):
                assert abs(r[key] - GOLD[pre + key][i]) < 1e-9


def test_K7_unit_profile_sample_equals_no_dla(oracle):
    params, model, samples, sp, prepared = prepared_config("tiny", 1, 8)
    nhi = samples.nhi_samples.copy(); nhi[2] = 0.0
    r = oracle.process_quasar_c(prepared[0], samples.offset_samples, nhi, -1.0, -1.0)
    assert r["sample_log_likelihoods_dla"][2] == r["log_likelihoods_no_dla"]


def test_extended_precision_restatement_is_pinned_by_mpmath():
    """oracle/lmlr_extended.c (log_mvnpdf_low_rank.m in x87 extended precision: the yardstick of tools/parity_truth.py)
    against a 50-digit dense evaluation of log N(y; mu, M M' + diag d): it must be at least two orders of magnitude
    closer to the exact value than the fp64 restatement's own rounding noise."""
    mp = pytest.importorskip("mpmath")
    from oracle import oracle as O
    mp.mp.dps = 50
    rng = np.random.default_rng(21)
    worst_ext, worst_f64 = 0.0, 0.0
    for n, k in ((24, 3), (60, 7), (120, 12)):
        y = rng.normal(size=n)
        mu = 0.1 * rng.normal(size=n)
        M = 0.5 * rng.normal(size=(n, k))
        d = rng.uniform(1e-3, 0.5, n)                      # low noise: the regime where q0 - t't cancels
        K = mp.matrix(n, n)
        for i in range(n):
            for j in range(n):
                K[i, j] = mp.fsum(mp.mpf(float(M[i, a])) * mp.mpf(float(M[j, a])) for a in range(k))
            K[i, i] += mp.mpf(float(d[i]))
        r = mp.matrix([mp.mpf(float(y[i])) - mp.mpf(float(mu[i])) for i in range(n)])
        quad = (r.T * mp.lu_solve(K, r))[0]
        exact = -(quad + mp.log(mp.det(K)) + n * mp.log(2 * mp.pi)) / 2
        ext = O.log_mvnpdf_low_rank_extended(y, mu, M, d)
        f64 = O.log_mvnpdf_low_rank_c(y, mu, M, d)
        worst_ext = max(worst_ext, abs(float(mp.mpf(ext) - exact) / float(exact)))
        worst_f64 = max(worst_f64, abs(float(mp.mpf(f64) - exact) / float(exact)))
    print(f"extended vs 50-digit exact: {worst_ext:.2e} relative; fp64 restatement: {worst_f64:.2e}")
    assert worst_ext <= 4e-16          # the final rounding to double
    assert worst_f64 <= 1e-11
