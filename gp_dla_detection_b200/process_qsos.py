"""Host-side mirror of the reference's ``process_qsos.m`` for the hot path.

The per-quasar preparation (process_qsos.m:102-120, :139-147, :160-180) is plain
numpy fp64 and follows the script statement by statement; the sample loop
(:184-198), the log-mean-exp (:202-212), the model posterior (:223-232) and the
MAP index (generate_ascii_catalog.m:73) run on the GPU through the C ABI of
``libdla_b200.so`` (include/dla_b200.h).  There is no CPU fallback: if the
library or a CUDA device is missing the calls raise.
"""
from __future__ import annotations

import ctypes as C
import logging
import time
from dataclasses import dataclass
from typing import Sequence

import numpy as np

from . import _lib
from .set_parameters import DEFAULT, Parameters


log = logging.getLogger("gp_dla_detection_b200")


@dataclass
class LearnedModel:
    """learned_qso_model_*.mat (learn_qso_model.m:113-123; loaded process_qsos.m:30-35)."""
    rest_wavelengths: np.ndarray  # [n_rest]
    mu: np.ndarray                # [n_rest]
    M: np.ndarray                 # [n_rest, k]
    log_omega: np.ndarray         # [n_rest]
    log_c_0: float
    log_tau_0: float
    log_beta: float


@dataclass
class DLASamples:
    """dla_samples.mat (generate_dla_samples.m:59-63; loaded process_qsos.m:37-40)."""
    offset_samples: np.ndarray   # [S]
    log_nhi_samples: np.ndarray  # [S]
    nhi_samples: np.ndarray      # [S]


@dataclass
class PreparedQuasar:
    z_qso: float
    n: int
    n_full: int
    padded_wavelengths: np.ndarray  # [n_full + 6]
    profile_index: np.ndarray       # [n] int32
    flux: np.ndarray
    mu: np.ndarray
    M: np.ndarray                   # [n, k] Fortran order (MATLAB memory)
    omega2: np.ndarray
    noise_variance: np.ndarray
    min_z_dla: float
    max_z_dla: float


def dla_existence_prior(z_qso, prior_z_qsos, prior_dla_ind, params: Parameters = DEFAULT):
    """process_qsos.m:123-132 -> (log_prior_dla, log_prior_no_dla); log(0) = -inf is legal."""
    less_ind = prior_z_qsos < (z_qso + params.prior_z_qso_increase)
    this_num_dlas = int(np.count_nonzero(np.asarray(prior_dla_ind)[less_ind]))
    this_num_quasars = int(np.count_nonzero(less_ind))
    with np.errstate(divide="ignore", invalid="ignore"):
        lp_dla = np.log(np.float64(this_num_dlas)) - np.log(np.float64(this_num_quasars))
        lp_no = np.log(np.float64(this_num_quasars - this_num_dlas)) - np.log(np.float64(this_num_quasars))
    return float(lp_dla), float(lp_no)


def _interp_columns(x, xp, fp):
    """griddedInterpolant(..., 'linear') on a 1-D grid (process_qsos.m:66-71), column-wise."""
    if fp.ndim == 1:
        return np.interp(x, xp, fp)
    out = np.empty((x.shape[0], fp.shape[1]), dtype=np.float64, order="F")
    for j in range(fp.shape[1]):
        out[:, j] = np.interp(x, xp, fp[:, j])
    return out


def prepare_quasar(wavelengths, flux, noise_variance, pixel_mask, z_qso: float, model: LearnedModel,
                   params: Parameters = DEFAULT, mask_gather: str = "reference") -> PreparedQuasar:
    """Steps 1-9 of SURVEY.md 3.1 (process_qsos.m:102-120, :139-147, :160-180).

    mask_gather="reference": the gather at :180/:189 as written, i.e. the first n
    elements of the n_full-long profile.  "intended": the unmasked positions.
    """
    if mask_gather not in ("reference", "intended"):
        raise ValueError("mask_gather must be 'reference' or 'intended'")
    this_wavelengths = np.asarray(wavelengths, dtype=np.float64)
    this_flux = np.asarray(flux, dtype=np.float64)
    this_noise_variance = np.asarray(noise_variance, dtype=np.float64)
    this_pixel_mask = np.asarray(pixel_mask, dtype=bool)

    this_rest_wavelengths = params.emitted_wavelengths(this_wavelengths, z_qso)            # :102
    ind = (this_rest_wavelengths >= params.min_lambda) & (this_rest_wavelengths <= params.max_lambda)  # :104-105
    this_unmasked_wavelengths = this_wavelengths[ind]                                       # :109
    in_range = ind
    ind = ind & (~this_pixel_mask)                                                          # :111
    this_wavelengths = this_wavelengths[ind]                                                # :113-116
    this_rest_wavelengths = this_rest_wavelengths[ind]
    this_flux = this_flux[ind]
    this_noise_variance = this_noise_variance[ind]
    n_full = int(this_unmasked_wavelengths.shape[0])
    n = int(this_wavelengths.shape[0])
    if n_full == 0 or n == 0:
        raise ValueError("quasar has no unmasked pixel in the modelled range")

    this_lya_zs = (this_wavelengths - params.lya_wavelength) / params.lya_wavelength        # :118-120

    this_mu = _interp_columns(this_rest_wavelengths, model.rest_wavelengths, model.mu)      # :139
    this_M = _interp_columns(this_rest_wavelengths, model.rest_wavelengths, model.M)        # :140
    this_log_omega = _interp_columns(this_rest_wavelengths, model.rest_wavelengths, model.log_omega)  # :142
    this_omega2 = np.exp(2 * this_log_omega)                                                # :143
    c_0, tau_0, beta = np.exp(model.log_c_0), np.exp(model.log_tau_0), np.exp(model.log_beta)  # :84-86
    this_scaling_factor = 1 - np.exp(-tau_0 * (1 + this_lya_zs) ** beta) + c_0              # :145
    this_omega2 = this_omega2 * this_scaling_factor ** 2                                    # :147

    min_z = params.min_z_dla(this_wavelengths, z_qso)                                       # :160
    max_z = params.max_z_dla(this_wavelengths, z_qso)                                       # :161

    w, ps = params.width, params.pixel_spacing                                              # :169-177
    lo = np.log10(np.min(this_unmasked_wavelengths))
    hi = np.log10(np.max(this_unmasked_wavelengths))
    padded = np.concatenate([
        10.0 ** np.linspace(lo - w * ps, lo - ps, w),
        this_unmasked_wavelengths,
        10.0 ** np.linspace(hi + ps, hi + w * ps, w),
    ])

    if mask_gather == "reference":
        # :180 `ind = ~this_pixel_mask(ind)` is all-true of length n, so :189 takes absorption(1:n)
        profile_index = np.arange(n, dtype=np.int32)
    else:
        profile_index = np.flatnonzero(~this_pixel_mask[in_range]).astype(np.int32)
    return PreparedQuasar(float(z_qso), n, n_full, np.ascontiguousarray(padded), profile_index,
                          np.ascontiguousarray(this_flux), np.ascontiguousarray(this_mu),
                          np.asfortranarray(this_M), np.ascontiguousarray(this_omega2),
                          np.ascontiguousarray(this_noise_variance), float(min_z), float(max_z))


class PackedBatch:
    """Ragged batch of prepared quasars in the CSR layout of dla_eval_batch (include/dla_b200.h)."""

    def __init__(self, quasars: Sequence[PreparedQuasar], k: int):
        Q = len(quasars)
        self.Q, self.k = Q, k
        ns = np.array([q.n for q in quasars], dtype=np.int64)
        Ps = np.array([q.n_full + 6 for q in quasars], dtype=np.int64)
        self.pix_ptr = np.concatenate([[0], np.cumsum(ns)]).astype(np.int64)
        self.pad_ptr = np.concatenate([[0], np.cumsum(Ps)]).astype(np.int64)
        cat = lambda xs, dt: (np.ascontiguousarray(np.concatenate(xs), dtype=dt) if Q else np.zeros(0, dt))
        self.padded_wavelengths = cat([q.padded_wavelengths for q in quasars], np.float64)
        self.profile_index = cat([q.profile_index for q in quasars], np.int32)
        self.flux = cat([q.flux for q in quasars], np.float64)
        self.mu = cat([q.mu for q in quasars], np.float64)
        self.omega2 = cat([q.omega2 for q in quasars], np.float64)
        self.noise_variance = cat([q.noise_variance for q in quasars], np.float64)
        for q in quasars:
            if q.M.shape != (q.n, k):
                raise ValueError("M must be n x k for every quasar")
        # per quasar column-major n x k, packed at k * pix_ptr[q]
        self.M = cat([np.asfortranarray(q.M).ravel(order="F") for q in quasars], np.float64)
        self.min_z_dla = np.array([q.min_z_dla for q in quasars], dtype=np.float64)
        self.max_z_dla = np.array([q.max_z_dla for q in quasars], dtype=np.float64)
        self.n = ns
        self.n_full = Ps - 6

    def input_bytes(self) -> int:
        return sum(a.nbytes for a in (self.pix_ptr, self.pad_ptr, self.padded_wavelengths, self.profile_index,
                                      self.flux, self.mu, self.omega2, self.noise_variance, self.M,
                                      self.min_z_dla, self.max_z_dla)) + 16 * self.Q


class PackedSpectra:
    """Ragged raw spectra (all_wavelengths / all_flux / all_noise_variance / all_pixel_mask of
    preloaded_qsos.mat, preload_qsos.m:73-79) packed for dla_eval_spectra."""

    def __init__(self, all_wavelengths, all_flux, all_noise_variance, all_pixel_mask, z_qsos):
        Q = len(all_wavelengths)
        self.Q = Q
        lens = np.array([len(w) for w in all_wavelengths], dtype=np.int64)
        self.spec_ptr = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        cat = lambda xs, dt: (np.ascontiguousarray(np.concatenate([np.asarray(x) for x in xs]), dtype=dt) if Q else np.zeros(0, dt))
        self.wavelengths = cat(all_wavelengths, np.float64)
        self.flux = cat(all_flux, np.float64)
        self.noise_variance = cat(all_noise_variance, np.float64)
        self.pixel_mask = cat(all_pixel_mask, np.uint8)
        self.z_qso = np.ascontiguousarray(z_qsos, dtype=np.float64)

    def input_bytes(self) -> int:
        return sum(a.nbytes for a in (self.spec_ptr, self.wavelengths, self.flux, self.noise_variance,
                                      self.pixel_mask, self.z_qso))


class DlaContext:
    """One ``dla_ctx`` (one CUDA device).  Thin, explicit wrapper of the C ABI."""

    def __init__(self, k: int = 20, num_lines: int = 3, num_samples: int = 10000, device: int = 0,
                 store_sample_ll: bool = True, workspace_bytes: int = 0):
        self._lib = _lib.load()
        self._ctx = C.c_void_p()
        self.k, self.num_lines, self.num_samples = int(k), int(num_lines), int(num_samples)
        self.store_sample_ll = bool(store_sample_ll)
        p = _lib.dla_params(self.k, self.num_lines, self.num_samples, int(self.store_sample_ll), 0,
                            int(workspace_bytes))
        rc = self._lib.dla_create(C.byref(self._ctx), int(device), C.byref(p))
        if rc != _lib.DLA_OK:
            msg = self._lib.dla_last_error(None).decode()
            self._ctx = C.c_void_p()
            raise _lib.DlaError(rc, msg)
        self._Q = 0
        self._keepalive = None

    # -- plumbing ---------------------------------------------------------------------------
    def _check(self, rc):
        if rc != _lib.DLA_OK:
            raise _lib.DlaError(rc, self._lib.dla_last_error(self._ctx).decode())

    def close(self):
        if getattr(self, "_ctx", None) is not None and self._ctx.value:
            self._lib.dla_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- API --------------------------------------------------------------------------------
    def set_samples(self, offset_samples, nhi_samples):
        off = np.ascontiguousarray(offset_samples, dtype=np.float64).ravel()
        nhi = np.ascontiguousarray(nhi_samples, dtype=np.float64).ravel()
        if off.shape[0] != self.num_samples or nhi.shape[0] != self.num_samples:
            raise ValueError("offset_samples / nhi_samples must have num_samples entries")
        self._check(self._lib.dla_set_samples(self._ctx, _lib.ptr(off, C.c_double), _lib.ptr(nhi, C.c_double)))

    def _input_args(self, b: PackedBatch, log_prior_dla, log_prior_no_dla):
        if b.k != self.k:
            raise ValueError("batch rank k differs from the context's")
        lpd = np.ascontiguousarray(np.broadcast_to(np.asarray(log_prior_dla, dtype=np.float64), (b.Q,)))
        lpn = np.ascontiguousarray(np.broadcast_to(np.asarray(log_prior_no_dla, dtype=np.float64), (b.Q,)))
        keep = (b, lpd, lpn)
        d, i64, i32 = C.c_double, C.c_int64, C.c_int32
        args = [b.Q, _lib.ptr(b.pix_ptr, i64), _lib.ptr(b.pad_ptr, i64), _lib.ptr(b.padded_wavelengths, d),
                _lib.ptr(b.profile_index, i32), _lib.ptr(b.flux, d), _lib.ptr(b.mu, d), _lib.ptr(b.omega2, d),
                _lib.ptr(b.noise_variance, d), _lib.ptr(b.M, d), _lib.ptr(b.min_z_dla, d), _lib.ptr(b.max_z_dla, d),
                _lib.ptr(lpd, d), _lib.ptr(lpn, d)]
        return args, keep

    def _alloc_outputs(self, Q, want_sample_ll=True, alloc=None):
        """Result arrays, NaN-initialised as process_qsos.m:74-82.  ``alloc(shape, dtype)`` may supply
        the memory (e.g. page-locked host buffers, so that the D2H copies run asynchronously)."""
        S = self.num_samples

        def full(shape, fill, dtype=np.float64):
            a = np.empty(shape, dtype=dtype) if alloc is None else alloc(shape, dtype)
            a[...] = fill
            return a
        out = {
            "sample_log_likelihoods_dla": (full((Q, S), np.nan) if (self.store_sample_ll and want_sample_ll) else None),
            "log_likelihoods_no_dla": full(Q, np.nan),
            "log_likelihoods_dla": full(Q, np.nan),
            "log_posteriors_no_dla": full(Q, np.nan),
            "log_posteriors_dla": full(Q, np.nan),
            "model_posteriors": full((Q, 2), np.nan),
            "map_ind": full(Q, -1, np.int32),
        }
        d = C.c_double
        args = [_lib.ptr(out["sample_log_likelihoods_dla"], d), _lib.ptr(out["log_likelihoods_no_dla"], d),
                _lib.ptr(out["log_likelihoods_dla"], d), _lib.ptr(out["log_posteriors_no_dla"], d),
                _lib.ptr(out["log_posteriors_dla"], d), _lib.ptr(out["model_posteriors"], d),
                _lib.ptr(out["map_ind"], C.c_int32)]
        return out, args

    def eval_batch(self, batch: PackedBatch, log_prior_dla, log_prior_no_dla, outputs=None):
        """dla_eval_batch: host buffers in, host buffers out (H2D and D2H inside the call)."""
        in_args, keep = self._input_args(batch, log_prior_dla, log_prior_no_dla)
        if outputs is None:
            outputs = self._alloc_outputs(batch.Q)
        out, out_args = outputs
        self._check(self._lib.dla_eval_batch(self._ctx, *in_args, *out_args))
        del keep
        return out

    def set_model(self, model: LearnedModel, params: Parameters = DEFAULT):
        """dla_set_model: the learned null model stays resident on the device (process_qsos.m:30-35, :84-86)."""
        if model.M.shape[1] != self.k:
            raise ValueError("model rank differs from the context's k")
        rest = np.ascontiguousarray(model.rest_wavelengths, dtype=np.float64)
        mu = np.ascontiguousarray(model.mu, dtype=np.float64)
        lo = np.ascontiguousarray(model.log_omega, dtype=np.float64)
        Mf = np.ascontiguousarray(np.asfortranarray(np.asarray(model.M, dtype=np.float64)).ravel(order="F"))
        pp = _lib.dla_prep_params(params.min_lambda, params.max_lambda, params.lya_wavelength, params.lyman_limit,
                                  params.max_z_cut, params.min_z_cut, params.pixel_spacing, int(params.width), 0,
                                  float(model.log_c_0), float(model.log_tau_0), float(model.log_beta))
        d = C.c_double
        self._check(self._lib.dla_set_model(self._ctx, rest.shape[0], _lib.ptr(rest, d), _lib.ptr(mu, d),
                                            _lib.ptr(Mf, d), _lib.ptr(lo, d), C.byref(pp)))

    def eval_spectra(self, spectra: "PackedSpectra", log_prior_dla, log_prior_no_dla, mask_gather: str = "reference",
                     outputs=None):
        """dla_eval_spectra: raw ragged spectra in (preloaded_qsos.mat arrays), results out; the per-quasar
        preparation of process_qsos.m:102-180 runs inside the library (selection on the host, model
        interpolation on the device)."""
        Q = spectra.Q
        lpd = np.ascontiguousarray(np.broadcast_to(np.asarray(log_prior_dla, dtype=np.float64), (Q,)))
        lpn = np.ascontiguousarray(np.broadcast_to(np.asarray(log_prior_no_dla, dtype=np.float64), (Q,)))
        if outputs is None:
            outputs = self._alloc_outputs(Q)
        out, out_args = outputs
        min_z, max_z = np.full(Q, np.nan), np.full(Q, np.nan)
        d = C.c_double
        self._check(self._lib.dla_eval_spectra(
            self._ctx, Q, _lib.ptr(spectra.spec_ptr, C.c_int64), _lib.ptr(spectra.wavelengths, d),
            _lib.ptr(spectra.flux, d), _lib.ptr(spectra.noise_variance, d), _lib.ptr(spectra.pixel_mask, C.c_uint8),
            _lib.ptr(spectra.z_qso, d), _lib.ptr(lpd, d), _lib.ptr(lpn, d),
            {"reference": 0, "intended": 1}[mask_gather], *out_args, _lib.ptr(min_z, d), _lib.ptr(max_z, d)))
        out = dict(out)
        out["min_z_dlas"], out["max_z_dlas"] = min_z, max_z
        return out

    def stage(self, batch: PackedBatch, log_prior_dla, log_prior_no_dla):
        in_args, keep = self._input_args(batch, log_prior_dla, log_prior_no_dla)
        self._check(self._lib.dla_stage_batch(self._ctx, *in_args))
        self._Q = batch.Q
        del keep

    def run_staged(self):
        self._check(self._lib.dla_run_staged(self._ctx))

    def fetch(self, want_sample_ll=True):
        out, out_args = self._alloc_outputs(self._Q, want_sample_ll)
        self._check(self._lib.dla_fetch_staged(self._ctx, *out_args))
        return out

    def timing(self) -> dict:
        t = _lib.dla_timing()
        self._check(self._lib.dla_get_timing(self._ctx, C.byref(t)))
        return t.as_dict()

    def voigt(self, lambdas, z, N, num_lines=None):
        """profile = voigt(lambdas, z, N, num_lines): the voigt MEX (voigt.c:253-304)."""
        lam = np.ascontiguousarray(lambdas, dtype=np.float64).ravel()
        out = np.empty(max(lam.shape[0] - 6, 0), dtype=np.float64)
        nl = 31 if num_lines is None else int(num_lines)  # voigt.c:16, :266
        self._check(self._lib.dla_voigt(self._ctx, _lib.ptr(lam, C.c_double), lam.shape[0], float(z), float(N), nl,
                                        _lib.ptr(out, C.c_double)))
        return out

    def voigt_noalloc(self, lambdas, z, N, num_lines=None):
        """dla_voigt_noalloc: voigt() on the library's static device scratch, no allocation in the call; at most
        _lib.NOALLOC_MAX_POINTS wavelengths (no counterpart in the reference; see include/dla_b200.h)."""
        lam = np.ascontiguousarray(lambdas, dtype=np.float64).ravel()
        out = np.empty(max(lam.shape[0] - 6, 0), dtype=np.float64)
        nl = 31 if num_lines is None else int(num_lines)
        self._check(self._lib.dla_voigt_noalloc(self._ctx, _lib.ptr(lam, C.c_double), lam.shape[0], float(z), float(N), nl,
                                                _lib.ptr(out, C.c_double)))
        return out

    def log_mvnpdf_low_rank(self, y, mu, M, d):
        """log_p = log_mvnpdf_low_rank(y, mu, M, d) (log_mvnpdf_low_rank.m:5)."""
        y = np.ascontiguousarray(y, dtype=np.float64).ravel()
        mu = np.ascontiguousarray(mu, dtype=np.float64).ravel()
        d = np.ascontiguousarray(d, dtype=np.float64).ravel()
        M = np.asarray(M, dtype=np.float64)
        n, k = M.shape
        Mf = np.ascontiguousarray(np.asfortranarray(M).ravel(order="F"))
        out = C.c_double()
        self._check(self._lib.dla_log_mvnpdf_low_rank(self._ctx, _lib.ptr(y, C.c_double), _lib.ptr(mu, C.c_double),
                                                      _lib.ptr(Mf, C.c_double), _lib.ptr(d, C.c_double), n, k,
                                                      C.byref(out)))
        return out.value


def finish_outputs(out: dict, batch: PackedBatch, samples: DLASamples, log_priors_dla, log_priors_no_dla) -> dict:
    """Add the remaining variables process_qsos.m saves (:235-243) and the MAP values of
    generate_ascii_catalog.m:73-80 to the arrays returned by the C ABI."""
    Q = batch.Q
    out = dict(out)
    out["min_z_dlas"] = batch.min_z_dla.copy()
    out["max_z_dlas"] = batch.max_z_dla.copy()
    out["log_priors_dla"] = np.broadcast_to(np.asarray(log_priors_dla, dtype=np.float64), (Q,)).copy()
    out["log_priors_no_dla"] = np.broadcast_to(np.asarray(log_priors_no_dla, dtype=np.float64), (Q,)).copy()
    out["p_no_dlas"] = out["model_posteriors"][:, 0].copy()          # :231
    out["p_dlas"] = 1 - out["p_no_dlas"]                              # :232
    mi = out["map_ind"]
    ok = mi >= 0
    map_z = np.full(Q, np.nan)
    map_nhi = np.full(Q, np.nan)
    off = samples.offset_samples
    # generate_ascii_catalog.m:75-80
    map_z[ok] = batch.min_z_dla[ok] + (batch.max_z_dla[ok] - batch.min_z_dla[ok]) * off[mi[ok]]
    map_nhi[ok] = samples.log_nhi_samples[mi[ok]]
    out["map_z_dlas"] = map_z
    out["map_log_nhis"] = map_nhi
    return out


def process_qsos(all_wavelengths, all_flux, all_noise_variance, all_pixel_mask, z_qsos, model: LearnedModel,
                 samples: DLASamples, log_priors_dla, log_priors_no_dla, params: Parameters = DEFAULT,
                 device: int = 0, mask_gather: str = "reference", store_sample_ll: bool = True,
                 context: DlaContext | None = None, device_prep: bool = False) -> dict:
    """Drop-in for the body of process_qsos.m:88-232 over a list of quasars.

    Inputs are the per-quasar ragged arrays of preloaded_qsos.mat (preload_qsos.m:73-79) and
    the per-quasar log priors (:129-132).  Returns a dict keyed by the variable names the
    script saves (:235-243) plus map_ind / map_z_dlas / map_log_nhis.
    """
    k = model.M.shape[1]
    S = samples.offset_samples.shape[0]
    t_start = time.perf_counter()
    log.info("processing %d quasars x %d DLA hypotheses (k = %d) on cuda:%d ...", len(z_qsos), S, k, device)

    def report(out):
        # process_qsos.m:93-94, :207-219 (fprintf / fprintf_debug), once per batch and per quasar at DEBUG level
        dt = time.perf_counter() - t_start
        log.info(" took %0.3fs (%.1f quasars/s).", dt, len(z_qsos) / max(dt, 1e-9))
        if log.isEnabledFor(logging.DEBUG):
            for i, z in enumerate(z_qsos):
                log.debug("quasar %i/%i (z_QSO = %0.4f): log p(D | z_QSO, no DLA) %0.2f, log p(D | z_QSO, DLA) %0.2f, "
                          "log p(DLA | D, z_QSO) %0.2f", i + 1, len(z_qsos), z, out["log_likelihoods_no_dla"][i],
                          out["log_likelihoods_dla"][i], out["log_posteriors_dla"][i])
        return out
    if device_prep:
        # process_qsos.m:102-180 inside the library (dla_set_model + dla_eval_spectra)
        spectra = PackedSpectra(all_wavelengths, all_flux, all_noise_variance, all_pixel_mask, z_qsos)
        own = context is None
        ctx = context or DlaContext(k=k, num_lines=params.num_lines, num_samples=S, device=device,
                                    store_sample_ll=store_sample_ll)
        try:
            if own:
                ctx.set_samples(samples.offset_samples, samples.nhi_samples)
            ctx.set_model(model, params)
            out = ctx.eval_spectra(spectra, log_priors_dla, log_priors_no_dla, mask_gather)
        finally:
            if own:
                ctx.close()

        class _Z:  # what finish_outputs reads from a PackedBatch
            Q, min_z_dla, max_z_dla = spectra.Q, out["min_z_dlas"], out["max_z_dlas"]
        return report(finish_outputs(out, _Z, samples, log_priors_dla, log_priors_no_dla))
    prepared = [prepare_quasar(w, f, v, m, z, model, params, mask_gather)
                for w, f, v, m, z in zip(all_wavelengths, all_flux, all_noise_variance, all_pixel_mask, z_qsos)]
    batch = PackedBatch(prepared, k)
    own = context is None
    ctx = context or DlaContext(k=k, num_lines=params.num_lines, num_samples=S, device=device,
                                store_sample_ll=store_sample_ll)
    try:
        if own:
            ctx.set_samples(samples.offset_samples, samples.nhi_samples)
        out = ctx.eval_batch(batch, log_priors_dla, log_priors_no_dla)
    finally:
        if own:
            ctx.close()
    return report(finish_outputs(out, batch, samples, log_priors_dla, log_priors_no_dla))
