"""-m gpu: the whole-catalogue MEX gateway (gp_dla_detection_b200/mex/dla_process_mex.c) EXECUTED: its mexFunction is
called with mxArrays built on the stand-in mex.h (no MATLAB / Octave on any box) and must return exactly what the Python
mirror returns for the same preloaded_qsos-style cell arrays (process_qsos.m:88-232, variables of :235-243)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from gp_dla_detection_b200 import _lib, synthetic
from gp_dla_detection_b200.process_qsos import PackedSpectra, process_qsos

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _build(tmp_path):
    pkg = os.path.join(ROOT, "gp_dla_detection_b200")
    _lib.load()
    out = tmp_path / "mex_process.so"
    env = dict(os.environ)
    env.pop("CC", None)
    subprocess.run(["gcc", "-O1", "-Wall", "-Werror", "-shared", "-fPIC", os.path.join(pkg, "mex", "dla_process_mex.c"),
                    os.path.join(ROOT, "tests", "mex_harness.c"), "-I", os.path.join(ROOT, "oracle", "shim"),
                    "-I", os.path.join(ROOT, "include"), "-L", pkg, "-l:libdla_b200.so", "-Wl,-rpath," + pkg,
                    "-Wl,--no-undefined", "-o", str(out)], check=True, env=env)
    return C.CDLL(str(out))


@pytest.mark.parametrize("mask_as_logical,mask_gather", [(1, "reference"), (0, "intended")])
def test_process_mex_equals_python_mirror(tmp_path, mask_as_logical, mask_gather):
    h = _build(tmp_path)
    Q, S, k = 9, 333, 20
    params, model, samples, sp = synthetic.make_config("config3", num_quasars=Q, num_samples=S)
    ref = process_qsos(sp.all_wavelengths, sp.all_flux, sp.all_noise_variance, sp.all_pixel_mask, sp.z_qsos, model, samples,
                       sp.log_priors_dla, sp.log_priors_no_dla, params, mask_gather=mask_gather, device_prep=True)
    ps = PackedSpectra(sp.all_wavelengths, sp.all_flux, sp.all_noise_variance, sp.all_pixel_mask, sp.z_qsos)
    par = np.array([params.min_lambda, params.max_lambda, params.lya_wavelength, params.lyman_limit,
                    params.max_z_cut, params.min_z_cut, params.pixel_spacing, params.width, params.num_lines,
                    1.0 if mask_gather == "intended" else 0.0])
    d = lambda a: np.ascontiguousarray(a, dtype=np.float64)
    outs = [np.full(Q, np.nan), np.full(Q, np.nan), np.full(Q, np.nan), np.full((S, Q), np.nan), np.full(Q, np.nan),
            np.full(Q, np.nan), np.full(Q, np.nan), np.full((2, Q), np.nan), np.full(Q, np.nan)]
    optr = (C.POINTER(C.c_double) * 9)(*[o.ctypes.data_as(C.POINTER(C.c_double)) for o in outs])
    P = C.POINTER(C.c_double)
    p = lambda a: a.ctypes.data_as(P)
    spec_ptr = np.ascontiguousarray(ps.spec_ptr, dtype=np.int64)
    wl, fl, nv, z = d(ps.wavelengths), d(ps.flux), d(ps.noise_variance), d(ps.z_qso)
    mask = np.ascontiguousarray(ps.pixel_mask, dtype=np.uint8)
    lpd, lpn = d(sp.log_priors_dla), d(sp.log_priors_no_dla)
    rest, mu, M, lo = d(model.rest_wavelengths), d(model.mu), np.asfortranarray(model.M, dtype=np.float64), d(model.log_omega)
    off, nhi = d(samples.offset_samples), d(samples.nhi_samples)
    h.harness_process.restype = C.c_int
    h.harness_last_error.restype = C.c_char_p
    rc = h.harness_process(C.c_long(Q), spec_ptr.ctypes.data_as(C.POINTER(C.c_long)), p(wl), p(fl), p(nv),
                           mask.ctypes.data_as(C.POINTER(C.c_ubyte)), C.c_int(mask_as_logical), p(z), p(lpd), p(lpn),
                           C.c_long(rest.shape[0]), C.c_int(k), p(rest), p(mu), M.ctypes.data_as(P), p(lo),
                           C.c_double(model.log_c_0), C.c_double(model.log_tau_0), C.c_double(model.log_beta),
                           C.c_long(S), p(off), p(nhi), p(par), C.c_int(9), optr)
    assert rc == 0, h.harness_last_error()
    h.harness_unload()
    names = ["min_z_dlas", "max_z_dlas", "log_likelihoods_no_dla", "sample_log_likelihoods_dla", "log_likelihoods_dla",
             "log_posteriors_no_dla", "log_posteriors_dla"]
    for name, got in zip(names, outs):
        want = np.asarray(ref[name])
        got = got.T if got.ndim == 2 else got            # MATLAB column-major Q x S  <->  numpy (S, Q)
        assert np.array_equal(got, want, equal_nan=True), name
    # model_posteriors(:, 1) is p_no_dlas (:231); p_dlas is 1 - p_no_dlas (:232), the second column the normalised exp (:229)
    assert np.array_equal(outs[7][0], ref["p_no_dlas"]) and np.array_equal(1 - outs[7][0], ref["p_dlas"])
    lp = np.stack([outs[5], outs[6]])
    want2 = np.exp(lp[1] - lp.max(axis=0)) / np.exp(lp - lp.max(axis=0)).sum(axis=0)
    assert np.allclose(outs[7][1], want2, rtol=1e-12, atol=0)
    assert np.array_equal(outs[8], np.asarray(ref["map_ind"]) + 1)


def test_unchanged_hypotheses_and_model_are_recognised_and_changed_ones_are_not():
    """dla_set_samples / dla_set_model return at once for the arrays already resident (the per-spectrum MEX loop passes
    them with every call); different values of the same length must still replace them."""
    from gp_dla_detection_b200.process_qsos import DlaContext
    Q, S = 3, 257
    params, model, samples, sp = synthetic.make_config("config3", num_quasars=Q, num_samples=S)
    ps = PackedSpectra(sp.all_wavelengths, sp.all_flux, sp.all_noise_variance, sp.all_pixel_mask, sp.z_qsos)
    with DlaContext(k=20, num_samples=S) as ctx:
        def run(off, nhi, mdl):
            ctx.set_samples(off, nhi)
            ctx.set_model(mdl, params)
            o = ctx.eval_spectra(ps, sp.log_priors_dla, sp.log_priors_no_dla)
            return {k: np.array(v, copy=True) for k, v in o.items()}
        a = run(samples.offset_samples, samples.nhi_samples, model)
        a2 = run(samples.offset_samples.copy(), samples.nhi_samples.copy(), model)     # same values, other addresses
        nhi_b = samples.nhi_samples.copy()
        nhi_b[5] *= 1.5
        b = run(samples.offset_samples, nhi_b, model)
        import dataclasses
        model_c = dataclasses.replace(model, mu=model.mu * 1.01)
        c = run(samples.offset_samples, nhi_b, model_c)
        a3 = run(samples.offset_samples, samples.nhi_samples, model)
    s = "sample_log_likelihoods_dla"
    assert np.array_equal(a[s], a2[s]) and np.array_equal(a[s], a3[s])
    assert not np.array_equal(a[s][:, 5], b[s][:, 5])
    keep = np.ones(S, bool); keep[5] = False
    assert np.array_equal(a[s][:, keep], b[s][:, keep])
    assert not np.array_equal(b["log_likelihoods_no_dla"], c["log_likelihoods_no_dla"])
    assert np.array_equal(a["log_likelihoods_no_dla"], a3["log_likelihoods_no_dla"])


def test_seam_entry_points_reuse_context_buffers_across_sizes():
    """dla_voigt / dla_log_mvnpdf_low_rank keep their device arrays in the context: growing, shrinking and repeated
    calls return what a fresh context returns."""
    from gp_dla_detection_b200.process_qsos import DlaContext
    rng = np.random.default_rng(7)
    def mvn(ctx, n, k):
        r = np.random.default_rng(n)
        y, mu, M, d = r.normal(size=n), r.normal(size=n) * 0.1, r.normal(size=(n, k)) * 0.3, r.uniform(0.1, 1.0, n)
        return ctx.log_mvnpdf_low_rank(y, mu, M, d)
    def voigt(ctx, n):
        lam = np.linspace(3700.0, 3700.0 + 0.9 * n, n) * 1e-8
        return ctx.voigt(lam, 2.3, 10 ** 20.5, 3)
    with DlaContext(k=20, num_samples=64) as ctx:
        got = [(mvn(ctx, n, 20), voigt(ctx, n)) for n in (300, 1500, 40, 1500, 300)]
    for (n, (gm, gv)) in zip((300, 1500, 40, 1500, 300), got):
        with DlaContext(k=20, num_samples=64) as fresh:
            assert mvn(fresh, n, 20) == gm
            assert np.array_equal(voigt(fresh, n), gv)
    del rng
