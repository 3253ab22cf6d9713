"""Seeded synthetic inputs for the hot path.

The reference ships no learned model, sample file, catalog or spectrum, so every
benchmark and parity input is generated here: numpy Generator(PCG64(seed)), fp64.
This is DATA generation only -- nothing in here evaluates the model.
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

from .process_qsos import DLASamples, LearnedModel
from .set_parameters import DEFAULT, Parameters


def halton(n: int, base: int) -> np.ndarray:
    """Points 1..n of the (unscrambled) van der Corput sequence in `base`."""
    idx = np.arange(1, n + 1, dtype=np.int64)
    out = np.zeros(n, dtype=np.float64)
    f = 1.0 / base
    while np.any(idx > 0):
        out += f * (idx % base)
        idx //= base
        f /= base
    return out


def make_model(k: int = 20, dlambda: float = 0.25, seed: int = 20260001, params: Parameters = DEFAULT) -> LearnedModel:
    rng = np.random.Generator(np.random.PCG64(seed))
    n_rest = int(round((params.max_lambda - params.min_lambda) / dlambda)) + 1
    rest = params.min_lambda + dlambda * np.arange(n_rest)  # learn_qso_model.m:29
    mu = (1 + 0.8 * np.exp(-((rest - 1215.67) / 12) ** 2) + 0.25 * np.exp(-((rest - 1033) / 8) ** 2)
          + 0.1 * (rest - 1060) / 150)
    phi = rng.uniform(0, 2 * np.pi, size=k)
    j = np.arange(k)
    s = 0.35 * (j + 1.0) ** -0.7
    span = params.max_lambda - params.min_lambda
    M = s[None, :] * np.cos(np.pi * (j[None, :] + 0.5) * (rest[:, None] - params.min_lambda) / span + phi[None, :])
    log_omega = np.log(0.12 + 0.08 * np.sin(np.pi * (rest - params.min_lambda) / 90) ** 2)
    return LearnedModel(rest, mu, np.asfortranarray(M), log_omega, float(np.log(params.initial_c_0)),
                        float(np.log(params.initial_tau_0)), float(np.log(params.initial_beta)))


def make_samples(num_samples: int, params: Parameters = DEFAULT) -> DLASamples:
    off = halton(num_samples, 2)
    log_nhi = params.uniform_min_log_nhi + (params.uniform_max_log_nhi - params.uniform_min_log_nhi) * halton(num_samples, 3)
    return DLASamples(off, log_nhi, 10.0 ** log_nhi)  # generate_dla_samples.m:57


def _voigt_profile_numpy(padded, z, N, num_lines=3):
    """Injected-DLA transmission for the synthetic flux (scipy wofz; generator only)."""
    from scipy.special import wofz
    c, sigma = 2.99792458e10, 9.08537121627923800e5
    tw = np.array([1.2156701e-05, 1.0257223e-05, 9.725368e-06])[:num_lines]
    lc = np.array([1.34347262962625339e-07, 2.15386482180851912e-08, 7.48525170087141461e-09])[:num_lines]
    gm = np.array([6.06075804241938613e+02, 1.54841462408931704e+02, 6.28964942715328164e+01])[:num_lines]
    ip = np.array([2.17460992138080811e-03, 4.11623059580451742e-02, 2.40309364651846963e-01,
                   4.32707438937454059e-01, 2.40309364651846963e-01, 4.11623059580451742e-02,
                   2.17460992138080811e-03])
    mult = c / (tw * (1 + z)) / 1e8
    v = padded[:, None] * mult[None, :] - c
    V = wofz((v + 1j * gm[None, :]) / (np.sqrt(2) * sigma)).real / (np.sqrt(2 * np.pi) * sigma)
    raw = np.exp(-N * (V * lc[None, :]).sum(axis=1))
    return np.convolve(raw, ip, mode="valid")


@dataclass
class SyntheticSpectra:
    """The ragged per-quasar arrays of preloaded_qsos.mat (preload_qsos.m:73-79)."""
    z_qsos: np.ndarray
    all_wavelengths: list
    all_flux: list
    all_noise_variance: list
    all_pixel_mask: list
    injected: np.ndarray  # bool: a DLA was injected
    log_priors_dla: np.ndarray
    log_priors_no_dla: np.ndarray


def make_spectra(num_quasars: int, model: LearnedModel, seed: int, ragged: bool = False,
                 params: Parameters = DEFAULT, inject: bool = True) -> SyntheticSpectra:
    """Configs 1/2/4/5 (ragged=False: n = n_full = n_rest, nothing masked) and config 3
    (ragged=True: blue cut at 3600 A observed + Poisson(3) masked runs, n >= min_num_pixels)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    from .process_qsos import prepare_quasar
    rest = model.rest_wavelengths
    k = model.M.shape[1]
    z_qsos = rng.uniform(params.z_qso_cut, 5.5, size=num_quasars)
    W, F, V, K = [], [], [], []
    injected = np.zeros(num_quasars, dtype=bool)
    for q in range(num_quasars):
        z = z_qsos[q]
        lam = rest * (1 + z)
        lam[0] *= (1 + 2.0 ** -40)   # keep the end points inside [min_lambda, max_lambda] after /(1+z)
        lam[-1] *= (1 - 2.0 ** -40)
        mask = np.zeros(lam.shape[0], dtype=bool)
        if ragged:
            lam = lam[lam >= 3600.0]
            while True:
                mask = np.zeros(lam.shape[0], dtype=bool)
                for _ in range(rng.poisson(3)):
                    ln = int(rng.integers(1, 26))
                    st = int(rng.integers(0, max(1, lam.shape[0] - ln)))
                    mask[st:st + ln] = True
                if np.count_nonzero(~mask) >= params.min_num_pixels:
                    break
        nv = np.clip(np.exp(np.log(0.05) + 0.7 * rng.standard_normal(lam.shape[0])), 1e-3, 4.0)
        flux0 = np.zeros(lam.shape[0])
        prep = prepare_quasar(lam, flux0, nv, np.zeros_like(mask), z, model, params)  # full grid, unmasked
        a = np.ones(prep.n)
        if inject and rng.random() < 0.5:
            injected[q] = True
            zs = rng.uniform(prep.min_z_dla, prep.max_z_dla)
            lognhi = rng.uniform(20.3, 22.0)
            a = _voigt_profile_numpy(prep.padded_wavelengths, zs, 10.0 ** lognhi, params.num_lines)
        zeta = rng.standard_normal(k)
        eps = rng.standard_normal(prep.n)
        flux = a * (prep.mu + prep.M @ zeta) + np.sqrt(prep.omega2 * a ** 2 + nv) * eps
        W.append(lam); F.append(flux); V.append(nv); K.append(mask)
    return SyntheticSpectra(z_qsos, W, F, V, K, injected,
                            np.full(num_quasars, np.log(0.1)), np.full(num_quasars, np.log(0.9)))


def make_config(name: str, num_quasars: int | None = None, num_samples: int | None = None):
    """Named workloads of BASELINE.json `configs`, optionally size-overridden.
    Returns (params, model, samples, spectra)."""
    table = {
        # name: (Q, S, k, dlambda, ragged, seed)
        "config1": (100, 10000, 20, 0.25, False, 20260001),
        "config2": (162861, 10000, 20, 0.25, False, 20260002),
        "config3": (2000, 10000, 20, 0.25, True, 20260003),
        "config4": (10000, 100000, 20, 0.25, False, 20260004),
        "config5": (20000, 10000, 40, 0.125, False, 20260005),
        "tiny": (4, 64, 5, 3.2, False, 20260000),
    }
    Q, S, k, dl, ragged, seed = table[name]
    Q = num_quasars if num_quasars is not None else Q
    S = num_samples if num_samples is not None else S
    params = Parameters(k=k, dlambda=dl, num_dla_samples=S)
    model = make_model(k, dl, seed, params)
    samples = make_samples(S, params)
    spectra = make_spectra(Q, model, seed + 1000, ragged, params)
    return params, model, samples, spectra


def make_training_set(num_quasars, k=20, n_rest=1217, seed=20260010, missing=0.15):
    """Synthetic stand-in for learn_qso_model.m:29-75: centred rest-frame fluxes on the model grid with NaN
    where a spectrum does not cover / is masked, (1 + z_Lya) per pixel, noise variances; and an x vector."""
    rng = np.random.default_rng(seed)
    rest = 911.75 + 0.25 * (1217 / n_rest) * np.arange(n_rest)
    z = rng.uniform(2.15, 5.0, num_quasars)
    l1pz = (rest[None, :] * (1 + z[:, None])) / 1215.6701
    M = np.stack([0.35 * (j + 1) ** -0.7 * np.cos(np.pi * (j + 0.5) * np.arange(n_rest) / n_rest + rng.uniform(0, 6.28))
                  for j in range(k)], axis=1)
    log_omega = np.log(0.12 + 0.08 * np.sin(np.pi * np.arange(n_rest) / 360) ** 2)
    nv = np.clip(np.exp(rng.normal(np.log(0.05), 0.7, (num_quasars, n_rest))), 1e-3, 4.0)
    c0, t0, b = 0.1, 0.0023, 3.65
    d = nv + np.exp(2 * log_omega)[None, :] * (1 - np.exp(-t0 * l1pz ** b) + c0) ** 2
    y = (rng.standard_normal((num_quasars, k)) @ M.T) + np.sqrt(d) * rng.standard_normal((num_quasars, n_rest))
    # missing data: a blue cut per quasar plus random masked runs (NaN in all three arrays, as :36-37, :57-60)
    for q in range(num_quasars):
        cut = int(rng.uniform(0, 0.4) * n_rest)
        bad = np.zeros(n_rest, bool)
        bad[:cut] = True
        bad |= rng.random(n_rest) < missing * rng.random()
        y[q, bad] = np.nan
        l1pz[q, bad] = np.nan
        nv[q, bad] = np.nan
    x = np.concatenate([(M * (1 + 0.05 * rng.standard_normal(M.shape))).ravel(order="F"),
                        log_omega + 0.05 * rng.standard_normal(n_rest), [np.log(0.11), np.log(0.0025), np.log(3.5)]])
    return x, y, l1pz, nv
