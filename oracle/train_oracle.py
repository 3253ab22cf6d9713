"""oracle/train_oracle.py -- TEST INFRASTRUCTURE ONLY.

numpy restatement, statement by statement, of the reference's training objective:
  spectrum_loss.m:14-76  negative log likelihood of one centred spectrum and its gradients
  objective.m:12-75      sum over the training set + the tau_0 / beta priors
Parity status: UNPINNED against MATLAB (no MATLAB/Octave here; the reference holds no tests or
goldens for these files).  What pins it instead: the gradient is checked against central finite
differences of the value, and the value against a dense n x n Gaussian (tests/test_train_oracle.py).
Only tests/ may import this module.
"""
from __future__ import annotations

import numpy as np
import scipy.linalg as sla

LOG_2PI = 1.83787706640934534  # spectrum_loss.m:17


def spectrum_loss(y, lya_1pz, noise_variance, M, omega2, c_0, tau_0, beta):
    n, k = M.shape                                                   # :19
    lya_optical_depth = tau_0 * lya_1pz ** beta                      # :22
    lya_absorption = np.exp(-lya_optical_depth)                      # :23
    scaling_factor = 1 - lya_absorption + c_0                        # :26
    absorption_noise = omega2 * scaling_factor ** 2                  # :27
    d = noise_variance + absorption_noise                            # :29
    d_inv = 1 / d                                                    # :31
    D_inv_y = d_inv * y                                              # :32
    D_inv_M = d_inv[:, None] * M                                     # :33
    B = M.T @ D_inv_M                                                # :40
    B[np.diag_indices(k)] += 1                                       # :41
    L = sla.cholesky(B, lower=False)                                 # :42  (MATLAB chol: upper, B = L'L)
    C = sla.solve_triangular(L, sla.solve_triangular(L, D_inv_M.T, trans="T"))  # :44  C = L \ (L' \ D_inv_M')
    K_inv_y = D_inv_y - D_inv_M @ (C @ y)                            # :46
    log_det_K = np.sum(np.log(d)) + 2 * np.sum(np.log(np.diag(L)))   # :48
    nlog_p = 0.5 * (y @ K_inv_y + log_det_K + n * LOG_2PI)           # :52
    K_inv_M = D_inv_M - D_inv_M @ (C @ M)                            # :55
    dM = -(np.outer(K_inv_y, K_inv_y @ M) - K_inv_M)                 # :56
    diag_K_inv = d_inv - np.sum(C * D_inv_M.T, axis=0)               # :59
    dlog_omega = -(absorption_noise * (K_inv_y ** 2 - diag_K_inv))   # :62
    da = c_0 * omega2 * scaling_factor                               # :65
    dlog_c_0 = -(K_inv_y * da) @ K_inv_y + diag_K_inv @ da           # :66
    da = omega2 * scaling_factor * lya_optical_depth * lya_absorption  # :69
    dlog_tau_0 = -(K_inv_y * da) @ K_inv_y + diag_K_inv @ da         # :70
    da = da * np.log(lya_1pz) * beta                                 # :73
    dlog_beta = -(K_inv_y * da) @ K_inv_y + diag_K_inv @ da          # :74
    return nlog_p, dM, dlog_omega, dlog_c_0, dlog_tau_0, dlog_beta


def objective(x, centered_rest_fluxes, lya_1pzs, rest_noise_variances, apply_priors=True):
    num_quasars, num_pixels = centered_rest_fluxes.shape             # :15
    k = (x.shape[0] - 3) // num_pixels - 1                           # :17
    M = x[:num_pixels * k].reshape((num_pixels, k), order="F")       # :19-20
    log_omega = x[num_pixels * k:num_pixels * (k + 1)]               # :22-23
    log_c_0, log_tau_0, log_beta = x[-3], x[-2], x[-1]               # :25-27
    omega2 = np.exp(2 * log_omega)                                   # :29
    c_0, tau_0, beta = np.exp(log_c_0), np.exp(log_tau_0), np.exp(log_beta)  # :30-32
    f = 0.0
    dM = np.zeros_like(M)
    dlog_omega = np.zeros_like(log_omega)
    dlog_c_0 = dlog_tau_0 = dlog_beta = 0.0
    for i in range(num_quasars):                                     # :41-57
        ind = ~np.isnan(centered_rest_fluxes[i])
        r = spectrum_loss(centered_rest_fluxes[i, ind], lya_1pzs[i, ind], rest_noise_variances[i, ind], M[ind],
                          omega2[ind], c_0, tau_0, beta)
        f += r[0]
        dM[ind] += r[1]
        dlog_omega[ind] += r[2]
        dlog_c_0 += r[3]
        dlog_tau_0 += r[4]
        dlog_beta += r[5]
    if apply_priors:
        dlog_tau_0 += tau_0 * (tau_0 - 0.0023) / 0.0007 ** 2          # :60-64  (Kim et al. 2007)
        dlog_beta += beta * (beta - 3.65) / 0.21 ** 2                 # :67-71
    g = np.concatenate([dM.ravel(order="F"), dlog_omega, [dlog_c_0, dlog_tau_0, dlog_beta]])  # :73
    return f, g


def dense_nlog_p(y, lya_1pz, noise_variance, M, omega2, c_0, tau_0, beta):
    """-log N(y; 0, M M' + diag(d)) by an n x n Cholesky (the definition at spectrum_loss.m:4-5)."""
    d = noise_variance + omega2 * (1 - np.exp(-tau_0 * lya_1pz ** beta) + c_0) ** 2
    K = M @ M.T + np.diag(d)
    Lc = np.linalg.cholesky(K)
    a = sla.solve_triangular(Lc, y, lower=True)
    return 0.5 * (a @ a + 2 * np.sum(np.log(np.diag(Lc))) + y.shape[0] * LOG_2PI)


from gp_dla_detection_b200.synthetic import make_training_set  # noqa: E402,F401  (synthetic inputs live in the package)
