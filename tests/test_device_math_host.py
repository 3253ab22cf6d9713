"""CPU check of the product's device Faddeeva formulation (csrc/faddeeva.cuh compiled with g++)
against the oracle's restatement of libcerf (cerf_port.c), scipy wofz and mpmath."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SIGMA = 9.08537121627923800e+05
GAMMAS = [6.06075804241938613e+02, 1.54841462408931704e+02, 6.28964942715328164e+01, 9.27433783998286437e-02]


@pytest.fixture(scope="module")
def host_math(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("hostmath") / "libdevice_math_host.so")
    env = dict(os.environ); env.pop("CC", None); env.pop("CXX", None)
    subprocess.run(["g++", "-O2", "-fPIC", "-shared", "-x", "c++", os.path.join(ROOT, "tests/host_shim/device_math_host.cpp"),
                    "-o", so], check=True, env=env)
    lib = C.CDLL(so)
    pd = C.POINTER(C.c_double)
    lib.host_re_w_many.argtypes = [pd, C.c_double, C.c_long, pd]

    def re_w(x, y):
        x = np.ascontiguousarray(x, dtype=np.float64)
        out = np.empty_like(x)
        lib.host_re_w_many(x.ctypes.data_as(pd), float(y), x.shape[0], out.ctypes.data_as(pd))
        return out
    return re_w


def _grid():
    rng = np.random.default_rng(7)
    return np.concatenate([np.linspace(0, 16, 4097), rng.uniform(0, 16, 50000), rng.uniform(15.9, 16.1, 2000),
                           rng.uniform(63, 65, 2000), 10 ** rng.uniform(np.log10(16), 4, 50000), -rng.uniform(0, 100, 1000)])


@pytest.mark.parametrize("gamma", GAMMAS)
def test_device_re_w_matches_wofz(host_math, gamma):
    from scipy.special import wofz
    y = gamma / np.sqrt(2.0) / SIGMA
    x = _grid()
    got = host_math(x, y)
    ref = wofz(x + 1j * y).real
    rel = np.abs(got - ref) / ref
    assert rel.max() < 5e-13, (rel.max(), x[rel.argmax()])


def test_device_re_w_matches_oracle_port(host_math, oracle):
    y = GAMMAS[0] / np.sqrt(2.0) / SIGMA
    x = np.concatenate([np.linspace(0, 20, 801), [50.0, 300.0, 3999.0, 4001.0, 7000.0]])
    got = host_math(x, y)
    ref = np.array([oracle.w_of_z_c(xx, y).real for xx in x])
    assert np.max(np.abs(got - ref) / ref) < 5e-13


def test_device_re_w_matches_mpmath(host_math):
    import mpmath as mp
    mp.mp.dps = 40
    for gamma in GAMMAS[:3]:
        y = gamma / np.sqrt(2.0) / SIGMA
        xs = np.array([0.0, 0.01, 0.7, 2.2, 3.9, 5.1, 6.0, 7.99, 8.01, 11.7, 15.99, 16.0, 16.01, 40.0, 64.0, 500.0, 6000.0])
        got = host_math(xs, y)
        for xx, g in zip(xs, got):
            z = mp.mpf(float(xx)) + 1j * mp.mpf(float(y))
            t = mp.re(mp.exp(-z * z) * mp.erfc(-1j * z))
            assert abs(g - t) / t < 5e-13, (xx, y)
