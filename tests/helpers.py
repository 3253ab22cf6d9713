"""Shared helpers for the parity tests."""
import numpy as np

from gp_dla_detection_b200 import synthetic
from gp_dla_detection_b200.process_qsos import PackedBatch, prepare_quasar

# north_star tolerances
RTOL_SAMPLE_LL = 1e-9      # per-sample log-likelihoods, relative
ATOL_LOG_POST = 1e-6       # model log-posteriors, absolute


def prepared_config(name, num_quasars, num_samples, mask_gather="reference"):
    params, model, samples, sp = synthetic.make_config(name, num_quasars=num_quasars, num_samples=num_samples)
    prepared = [prepare_quasar(w, f, v, m, z, model, params, mask_gather)
                for w, f, v, m, z in zip(sp.all_wavelengths, sp.all_flux, sp.all_noise_variance,
                                         sp.all_pixel_mask, sp.z_qsos)]
    return params, model, samples, sp, prepared


def oracle_batch(O, prepared, samples, lp_dla, lp_no, num_lines=3, use_ref_voigt=None):
    """Run the C oracle over a list of prepared quasars; returns dict of stacked arrays."""
    if use_ref_voigt is None:
        use_ref_voigt = O.load_ref() is not None
    rows = [O.process_quasar_c(p, samples.offset_samples, samples.nhi_samples, lp_dla[i], lp_no[i], num_lines,
                               use_ref_voigt=use_ref_voigt) for i, p in enumerate(prepared)]
    out = {k: np.stack([np.asarray(r[k]) for r in rows]) for k in rows[0]}
    return out


def assert_parity(got, ref, check_map=True):
    """The north_star gates: 1e-9 relative per-sample ll, 1e-6 absolute log posteriors,
    identical MAP index (when the top-2 gap is resolvable) and identical DLA decision."""
    a, b = got["sample_log_likelihoods_dla"], ref["sample_log_likelihoods_dla"]
    assert a.shape == b.shape
    rel = np.abs(a - b) / np.maximum(np.abs(b), 1e-300)
    assert np.nanmax(rel) <= RTOL_SAMPLE_LL, f"per-sample ll rel err {np.nanmax(rel):.3e}"
    for key in ("log_likelihoods_no_dla", "log_likelihoods_dla", "log_posteriors_no_dla", "log_posteriors_dla"):
        same = got[key] == ref[key]          # covers -inf priors (process_qsos.m:129-132)
        with np.errstate(invalid="ignore"):
            err = np.max(np.where(same, 0.0, np.abs(got[key] - ref[key])))
        assert err <= ATOL_LOG_POST, f"{key} abs err {err:.3e}"
    assert np.max(np.abs(got["model_posteriors"] - ref["model_posteriors"])) <= 1e-6
    dec_got = got["log_posteriors_dla"] > got["log_posteriors_no_dla"]
    dec_ref = ref["log_posteriors_dla"] > ref["log_posteriors_no_dla"]
    margin = np.abs(ref["log_posteriors_dla"] - ref["log_posteriors_no_dla"])
    assert np.all((dec_got == dec_ref) | (margin < 10 * ATOL_LOG_POST))
    if check_map:
        for q in range(b.shape[0]):
            srt = np.sort(b[q][~np.isnan(b[q])])
            gap = srt[-1] - srt[-2] if srt.shape[0] > 1 else np.inf
            if gap > 10 * RTOL_SAMPLE_LL * abs(srt[-1]):   # only well posed if resolvable
                assert got["map_ind"][q] == ref["map_ind"][q], f"MAP index differs for quasar {q} (gap {gap:.3e})"
    return float(np.nanmax(rel))
