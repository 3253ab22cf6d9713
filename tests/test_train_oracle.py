"""The training-objective oracle (spectrum_loss.m / objective.m restated in numpy): value against the
dense Gaussian it is defined as, gradient against central finite differences."""
import numpy as np

from oracle import train_oracle as T


def test_spectrum_loss_value_equals_dense_gaussian():
    x, y, l1, nv = T.make_training_set(3, k=5, n_rest=90, seed=1)
    n, k = 90, 5
    M = x[:n * k].reshape((n, k), order="F")
    om2 = np.exp(2 * x[n * k:n * (k + 1)])
    c0, t0, b = np.exp(x[-3:])
    for q in range(3):
        ind = ~np.isnan(y[q])
        v = T.spectrum_loss(y[q, ind], l1[q, ind], nv[q, ind], M[ind], om2[ind], c0, t0, b)[0]
        ref = T.dense_nlog_p(y[q, ind], l1[q, ind], nv[q, ind], M[ind], om2[ind], c0, t0, b)
        assert abs(v - ref) <= 1e-10 * abs(ref)


def test_objective_gradient_matches_finite_differences():
    x, y, l1, nv = T.make_training_set(4, k=3, n_rest=40, seed=2)
    f, g = T.objective(x, y, l1, nv, apply_priors=False)
    rng = np.random.default_rng(0)
    idx = np.concatenate([rng.choice(x.shape[0] - 3, 25, replace=False), [x.shape[0] - 3, x.shape[0] - 2, x.shape[0] - 1]])
    for i in idx:
        h = 1e-5 * max(1.0, abs(x[i]))
        xp, xm = x.copy(), x.copy()
        xp[i] += h
        xm[i] -= h
        fd = (T.objective(xp, y, l1, nv, False)[0] - T.objective(xm, y, l1, nv, False)[0]) / (2 * h)
        assert abs(fd - g[i]) <= 2e-6 * max(1.0, abs(g[i])), (i, fd, g[i])
    # the priors only touch the last two entries (objective.m:60-71)
    g2 = T.objective(x, y, l1, nv, True)[1]
    assert np.array_equal(g2[:-2], g[:-2]) and g2[-2] != g[-2] and g2[-1] != g[-1]
