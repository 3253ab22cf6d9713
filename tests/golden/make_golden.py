#!/usr/bin/env python
"""Generate the committed golden vectors (run in the dev container, where /root/reference exists).

The reference is MATLAB + a C MEX file and ships no tests or fixtures, so the goldens are
produced by the parts of the reference that DO run here plus the restatements pinned to them:
  voigt_profiles   the reference's voigt.c compiled VERBATIM (oracle/_ref/libvoigt_ref.so; libcerf
                   replaced by oracle/cerf_port.c)                         -> pins voigt.c:278-299
  faddeeva         scipy.special.wofz (the Faddeeva package libcerf 1.x wraps) -> pins cerf_port.c
  tiny / config1_subset  numpy restatement (O2n) resp. C restatement with voigt.c verbatim (O2c+O1)
                   of process_qsos.m:150-212, :223-229 + log_mvnpdf_low_rank.m as written
Inputs are regenerated in the tests from the seeded generator; an input checksum guards drift.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import oracle as O  # noqa: E402
from tests.helpers import prepared_config  # noqa: E402

VOIGT_CASES = [(0.5, 0.0), (0.1, 1e20), (0.5, 10 ** 20.5), (0.9, 1e21), (0.3, 1e22), (0.7, 1e23)]


def checksum(prepared):
    h = hashlib.sha256()
    for p in prepared:
        for a in (p.padded_wavelengths, p.flux, p.mu, np.ascontiguousarray(np.asarray(p.M)), p.omega2, p.noise_variance):
            h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def main():
    assert O.load_ref() is not None, "needs /root/reference/voigt.c"
    from scipy.special import wofz
    out = {}
    # 1. voigt.c verbatim
    params, model, samples, sp, prepared = prepared_config("config1", 2, 8)
    p = prepared[0]
    out["voigt_lambdas"] = p.padded_wavelengths
    zs = np.array([p.min_z_dla + (p.max_z_dla - p.min_z_dla) * f for f, _ in VOIGT_CASES])
    Ns = np.array([n for _, n in VOIGT_CASES])
    out["voigt_z"], out["voigt_N"] = zs, Ns
    out["voigt_profiles"] = np.stack([O.voigt_ref(p.padded_wavelengths, z, n, 3) for z, n in zip(zs, Ns)])
    out["voigt_profile_31"] = O.voigt_ref(p.padded_wavelengths, zs[3], 1e21, 31)
    # 2. Faddeeva
    rng = np.random.default_rng(11)
    x = np.concatenate([np.linspace(0, 20, 401), 10 ** rng.uniform(0, 3.9, 400)])
    ys = np.array([g / np.sqrt(2.0) / O.SIGMA for g in O.GAMMAS])
    out["fadd_x"], out["fadd_y"] = x, ys
    out["fadd_re_w"] = np.stack([wofz(x + 1j * y).real for y in ys])
    # 3. tiny config through the numpy restatement
    params, model, samples, sp, prepared = prepared_config("tiny", 4, 64)
    rows = [O.process_quasar_numpy(q, samples.offset_samples, samples.nhi_samples, sp.log_priors_dla[i],
                                   sp.log_priors_no_dla[i]) for i, q in enumerate(prepared)]
    out["tiny_checksum"] = np.array(checksum(prepared))
    for key in rows[0]:
        out["tiny_" + key] = np.stack([np.asarray(r[key]) for r in rows])
    # 4. config-1 shape subset through the C restatement with voigt.c verbatim
    params, model, samples, sp, prepared = prepared_config("config1", 2, 96)
    rows = [O.process_quasar_c(q, samples.offset_samples, samples.nhi_samples, sp.log_priors_dla[i],
                               sp.log_priors_no_dla[i], use_ref_voigt=True) for i, q in enumerate(prepared)]
    out["c1_checksum"] = np.array(checksum(prepared))
    for key in rows[0]:
        out["c1_" + key] = np.stack([np.asarray(r[key]) for r in rows])
    np.savez_compressed(os.path.join(HERE, "golden_v1.npz"), **out)
    print("wrote golden_v1.npz", {k: np.asarray(v).shape for k, v in out.items()})


if __name__ == "__main__":
    main()
