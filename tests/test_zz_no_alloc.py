"""-m gpu, collected LAST (a failure here must not hide another test).

dla_voigt_noalloc -- the entry point for callers that cannot call cudaMalloc (no counterpart in the reference):
the voigt seam on the library's file-scope static device scratch.  Through the
C ABI, against dla_voigt (the existing entry point, same kernel) on two calls in a row whose sizes differ, so that
the second call runs on the first call's leftovers; sizes beyond the static maximum are refused with a clear error.
Passes alone in a fresh process and after the other tests (it creates its own context and relies on no state).
"""
import ctypes as C

import numpy as np
import pytest

from gp_dla_detection_b200 import _lib
from gp_dla_detection_b200.process_qsos import DlaContext
from tests.helpers import RTOL_SAMPLE_LL

pytestmark = pytest.mark.gpu


def _grid(n, lo=3650.0):
    return lo * 10.0 ** (1e-4 * np.arange(n))          # log-spaced observed wavelengths, Angstrom (voigt.c takes Angstrom)


def test_noalloc_entry_point_matches_dla_voigt_on_two_calls_in_a_row():
    calls = [(_grid(6000), 2.31, 10 ** 20.9, 3),         # first: a long grid fills 6000 entries of each scratch part
             (_grid(1237, 3900.0), 2.55, 10 ** 21.4, 3),  # second: a short one, on the first call's leftovers
             (_grid(1237, 3900.0), 2.55, 10 ** 21.4, 31)]
    with DlaContext(k=20, num_samples=64) as ctx:
        for lam, z, N, nl in calls:
            want = ctx.voigt(lam, z, N, nl)
            got = ctx.voigt_noalloc(lam, z, N, nl)
            assert got.shape == want.shape == (lam.shape[0] - 6,)
            assert np.all(np.isfinite(got)) and got.min() < 0.5 < got.max()      # a damped line is in the grid
            assert np.allclose(got, want, rtol=RTOL_SAMPLE_LL, atol=0.0)
            assert np.array_equal(got, want)                                     # same kernel, same inputs
        # a second context of the same device shares the array and its address
        with DlaContext(k=20, num_samples=64) as other:
            lam, z, N, nl = calls[1]
            assert np.array_equal(other.voigt_noalloc(lam, z, N, nl), ctx.voigt(lam, z, N, nl))


def test_noalloc_entry_point_rejects_sizes_beyond_its_static_scratch():
    lam = _grid(_lib.NOALLOC_MAX_POINTS + 1)
    out = np.full(lam.shape[0] - 6, -7.0)
    with DlaContext(k=20, num_samples=64) as ctx:
        rc = ctx._lib.dla_voigt_noalloc(ctx._ctx, _lib.ptr(lam, C.c_double), lam.shape[0], 2.3, 1e21, 3, _lib.ptr(out, C.c_double))
        assert rc == -1                                                          # DLA_ERR_BAD_ARG
        assert b"DLA_NOALLOC_MAX_POINTS" in ctx._lib.dla_last_error(ctx._ctx)
        assert np.all(out == -7.0)                                               # nothing was written
        with pytest.raises(_lib.DlaError, match="DLA_NOALLOC_MAX_POINTS"):
            ctx.voigt_noalloc(lam, 2.3, 1e21, 3)
        with pytest.raises(_lib.DlaError, match="at least 7"):
            ctx.voigt_noalloc(lam[:6], 2.3, 1e21, 3)
        full = _grid(_lib.NOALLOC_MAX_POINTS)                                    # the maximum itself is served
        assert np.array_equal(ctx.voigt_noalloc(full, 2.3, 1e21, 3), ctx.voigt(full, 2.3, 1e21, 3))
