"""Training side of the pipeline on the same library.

  objective.m:12-75 / spectrum_loss.m:14-76  -> TrainingObjective (CUDA, through the C ABI)
  learn_qso_model.m:29-75                    -> prepare_training_set (numpy, statement by statement)
  learn_qso_model.m:77-111                   -> learn_qso_model (pairwise-PCA initialisation + L-BFGS)

The objective is data-parallel over training quasars: every rank holds a slice of the spectra, evaluates
its partial sum on its own GPU and the partial (f, g) are SUMMED over ranks -- the one all-reduce of the
pipeline (``dist`` = an initialised ``torch.distributed``; NCCL on GPUs, gloo in the CPU tests).  The tau_0
and beta priors (objective.m:60-71) are added once, after the reduction.  minFunc (the reference's L-BFGS)
is an external MATLAB package; scipy's L-BFGS-B drives the same objective here.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from .process_qsos import DlaContext, LearnedModel
from .set_parameters import DEFAULT, Parameters

TAU_0_MU, TAU_0_SIGMA = 0.0023, 0.0007   # objective.m:60-61 (Kim et al. 2007)
BETA_MU, BETA_SIGMA = 3.65, 0.21         # objective.m:67-68


def add_priors(x, g):
    """objective.m:60-71: the priors touch the gradient only (as written in the reference)."""
    tau_0, beta = np.exp(x[-2]), np.exp(x[-1])
    g = g.copy()
    g[-2] += tau_0 * (tau_0 - TAU_0_MU) / TAU_0_SIGMA ** 2
    g[-1] += beta * (beta - BETA_MU) / BETA_SIGMA ** 2
    return g


def reduce_over_ranks(f, g, dist=None):
    """Sum the per-rank partial objective and gradient (the pipeline's only collective)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return f, g
    import torch
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.from_numpy(np.concatenate([[f], g])).to(dev)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    t = t.cpu().numpy()
    return float(t[0]), t[1:].copy()


class TrainingObjective:
    """f, g = obj(x) with the training slice resident on the device (dla_train_set_data / dla_train_objective)."""

    def __init__(self, centered_rest_fluxes, lya_1pzs, rest_noise_variances, k: int, device: int = 0, dist=None,
                 context: DlaContext | None = None):
        y = np.ascontiguousarray(centered_rest_fluxes, dtype=np.float64)
        l1 = np.ascontiguousarray(lya_1pzs, dtype=np.float64)
        nv = np.ascontiguousarray(rest_noise_variances, dtype=np.float64)
        if y.ndim != 2 or l1.shape != y.shape or nv.shape != y.shape:
            raise ValueError("centered_rest_fluxes, lya_1pzs, rest_noise_variances must be [num_quasars, num_pixels]")
        self.num_quasars, self.num_pixels = y.shape
        self.k, self.dist = int(k), dist
        self._own = context is None
        self.ctx = context or DlaContext(k=self.k, num_samples=0, device=device)
        d = C.c_double
        self.ctx._check(self.ctx._lib.dla_train_set_data(self.ctx._ctx, self.num_quasars, self.num_pixels,
                                                         _lib.ptr(y, d), _lib.ptr(l1, d), _lib.ptr(nv, d)))
        self.nx = self.num_pixels * (self.k + 1) + 3

    def local(self, x):
        """This rank's partial sums, no priors."""
        x = np.ascontiguousarray(x, dtype=np.float64).ravel()
        if x.shape[0] != self.nx:
            raise ValueError(f"x must have num_pixels*(k+1)+3 = {self.nx} entries")
        f = C.c_double()
        g = np.empty(self.nx)
        self.ctx._check(self.ctx._lib.dla_train_objective(self.ctx._ctx, _lib.ptr(x, C.c_double), 0, C.byref(f),
                                                          _lib.ptr(g, C.c_double)))
        return f.value, g

    def __call__(self, x, apply_priors: bool = True):
        f, g = reduce_over_ranks(*self.local(x), self.dist)
        return f, (add_priors(np.asarray(x, dtype=np.float64).ravel(), g) if apply_priors else g)

    def close(self):
        if self._own:
            self.ctx.close()


def objective(x, centered_rest_fluxes, lya_1pzs, rest_noise_variances, device: int = 0):
    """[f, g] = objective(x, centered_rest_fluxes, lya_1pzs, rest_noise_variances)  (objective.m:12)."""
    num_pixels = np.asarray(centered_rest_fluxes).shape[1]
    k = (np.asarray(x).shape[0] - 3) // num_pixels - 1          # objective.m:17
    obj = TrainingObjective(centered_rest_fluxes, lya_1pzs, rest_noise_variances, k, device)
    try:
        return obj(x)
    finally:
        obj.close()


def _interp1(xq, x, v):
    """interp1(x, v, xq): linear, NaN outside the data range (MATLAB default)."""
    return np.interp(xq, x, v, left=np.nan, right=np.nan)


def prepare_training_set(all_wavelengths, all_flux, all_noise_variance, all_pixel_mask, z_qsos,
                         params: Parameters = DEFAULT, max_noise_variance: float = 1.0 ** 2):
    """learn_qso_model.m:29-75 -> rest_wavelengths, mu, centered_rest_fluxes, lya_1pzs, rest_noise_variances."""
    rest_wavelengths = np.arange(params.min_lambda, params.max_lambda + params.dlambda / 2, params.dlambda)  # :29
    Q, n = len(z_qsos), rest_wavelengths.shape[0]
    lya_1pzs = np.full((Q, n), np.nan)                                                  # :32-34
    rest_fluxes = np.full((Q, n), np.nan)
    rest_noise_variances = np.full((Q, n), np.nan)
    for i in range(Q):                                                                   # :37-60
        w = np.asarray(all_wavelengths[i], dtype=np.float64)
        fl = np.asarray(all_flux[i], dtype=np.float64).copy()
        nv = np.asarray(all_noise_variance[i], dtype=np.float64).copy()
        m = np.asarray(all_pixel_mask[i], dtype=bool)
        fl[m] = np.nan                                                                   # :45-46
        nv[m] = np.nan
        rest = params.emitted_wavelengths(w, z_qsos[i])                                  # :48
        lya_1pzs[i] = _interp1(rest_wavelengths, rest, 1 + (w - params.lya_wavelength) / params.lya_wavelength)  # :50-53
        rest_fluxes[i] = _interp1(rest_wavelengths, rest, fl)                            # :55-56
        rest_noise_variances[i] = _interp1(rest_wavelengths, rest, nv)                   # :58-59
    ind = rest_noise_variances > max_noise_variance                                      # :64-67
    lya_1pzs[ind] = np.nan
    rest_fluxes[ind] = np.nan
    rest_noise_variances[ind] = np.nan
    # a pixel is usable only where all three are defined (objective.m:42 tests the flux alone)
    bad = np.isnan(rest_fluxes) | np.isnan(lya_1pzs) | np.isnan(rest_noise_variances)
    rest_fluxes[bad] = np.nan
    with np.errstate(invalid="ignore"):
        mu = np.nanmean(rest_fluxes, axis=0)                                             # :70
    centered = rest_fluxes - mu[None, :]                                                 # :71
    return rest_wavelengths, mu, centered, lya_1pzs, rest_noise_variances


def initial_x(centered_rest_fluxes, k: int, params: Parameters = DEFAULT):
    """learn_qso_model.m:74-96: top-k pairwise-complete PCA for M, log nanstd for log omega."""
    y = np.asarray(centered_rest_fluxes, dtype=np.float64)
    ok = ~np.isnan(y)
    y0 = np.where(ok, y, 0.0)
    cnt = ok.T.astype(np.float64) @ ok.astype(np.float64)
    s = y0.T @ y0
    m1 = y0.T @ ok.astype(np.float64)           # sum of x_i over the rows where both are present
    with np.errstate(invalid="ignore", divide="ignore"):
        cov = (s - m1 * m1.T / cnt) / (cnt - 1)  # pca(..., 'rows', 'pairwise')
    cov = np.where(np.isfinite(cov), cov, 0.0)
    cov = 0.5 * (cov + cov.T)
    lat, vec = np.linalg.eigh(cov)
    order = np.argsort(lat)[::-1][:k]
    latent, coeff = np.maximum(lat[order], 0.0), vec[:, order]
    M0 = coeff * np.sqrt(latent)[None, :]                                                # :85
    with np.errstate(invalid="ignore"):
        lo0 = np.log(np.nanstd(y, axis=0, ddof=1))                                       # :88
    lo0 = np.where(np.isfinite(lo0), lo0, np.log(0.1))
    return np.concatenate([M0.ravel(order="F"), lo0, [np.log(params.initial_c_0), np.log(params.initial_tau_0),
                                                     np.log(params.initial_beta)]])     # :90-98


def learn_qso_model(rest_wavelengths, mu, centered_rest_fluxes, lya_1pzs, rest_noise_variances, k: int = 20,
                    params: Parameters = DEFAULT, device: int = 0, dist=None, max_iter: int = 2000, max_fun: int = 4000,
                    x0=None):
    """learn_qso_model.m:77-111 with scipy's L-BFGS-B in place of minFunc.  Returns (LearnedModel, f, info)."""
    from scipy.optimize import minimize
    obj = TrainingObjective(centered_rest_fluxes, lya_1pzs, rest_noise_variances, k, device, dist)
    try:
        x0 = initial_x(centered_rest_fluxes, k, params) if x0 is None else np.asarray(x0, dtype=np.float64)
        res = minimize(lambda x: obj(x), x0, jac=True, method="L-BFGS-B", options={"maxiter": max_iter, "maxfun": max_fun, "maxcor": 100})
    finally:
        obj.close()
    n = np.asarray(rest_wavelengths).shape[0]
    x = res.x
    M = x[:n * k].reshape((n, k), order="F")                                             # :103-104
    model = LearnedModel(np.asarray(rest_wavelengths, dtype=np.float64), np.asarray(mu, dtype=np.float64),
                         np.asfortranarray(M), x[n * k:n * (k + 1)].copy(), float(x[-3]), float(x[-2]), float(x[-1]))
    return model, float(res.fun), res
