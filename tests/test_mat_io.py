"""Round trips of the reference's on-disk variables through MATLAB v5/v7 files (scipy.io); the v7.3/HDF5
branch needs h5py (absent here) and must say so instead of mis-reading."""
import numpy as np
import pytest
from scipy.io import savemat

from gp_dla_detection_b200 import mat_io, synthetic


def test_model_samples_and_spectra_load_by_reference_variable_names(tmp_path):
    params, model, samples, sp = synthetic.make_config("tiny", num_quasars=3, num_samples=16)
    savemat(tmp_path / "learned_qso_model_x.mat",
            {"rest_wavelengths": model.rest_wavelengths, "mu": model.mu, "M": model.M, "log_omega": model.log_omega,
             "log_c_0": model.log_c_0, "log_tau_0": model.log_tau_0, "log_beta": model.log_beta}, oned_as="column")
    m2 = mat_io.load_learned_model(str(tmp_path / "learned_qso_model_x.mat"))
    assert np.array_equal(m2.M, model.M) and np.array_equal(m2.mu, model.mu) and m2.log_beta == model.log_beta
    savemat(tmp_path / "dla_samples.mat", {"offset_samples": samples.offset_samples[None, :],      # 1 x S rows
                                           "log_nhi_samples": samples.log_nhi_samples[None, :],
                                           "nhi_samples": samples.nhi_samples[None, :]})
    s2 = mat_io.load_dla_samples(str(tmp_path / "dla_samples.mat"))
    assert np.array_equal(s2.nhi_samples, samples.nhi_samples)

    def cell(xs):
        c = np.empty((len(xs), 1), dtype=object)
        for i, x in enumerate(xs):
            c[i, 0] = np.asarray(x, dtype=np.float64).reshape(-1, 1)
        return c
    savemat(tmp_path / "preloaded_qsos.mat", {"all_wavelengths": cell(sp.all_wavelengths), "all_flux": cell(sp.all_flux),
                                              "all_noise_variance": cell(sp.all_noise_variance),
                                              "all_pixel_mask": cell(sp.all_pixel_mask)})
    p = mat_io.load_preloaded_qsos(str(tmp_path / "preloaded_qsos.mat"))
    for q in range(3):
        assert np.array_equal(p["all_wavelengths"][q], sp.all_wavelengths[q])
        assert np.array_equal(p["all_pixel_mask"][q], sp.all_pixel_mask[q]) and p["all_pixel_mask"][q].dtype == bool


def test_processed_qsos_round_trip(tmp_path):
    Q, S = 4, 9
    rng = np.random.default_rng(1)
    res = {n: rng.standard_normal(Q) for n in mat_io.RESULT_VARIABLES}
    res["sample_log_likelihoods_dla"] = rng.standard_normal((Q, S))
    res["sample_log_likelihoods_dla"][1, 3] = np.nan
    res["model_posteriors"] = rng.random((Q, 2))
    res["log_priors_dla"][0] = -np.inf
    path = str(tmp_path / "processed_qsos_t.mat")
    mat_io.save_processed_qsos(path, res, extra={"num_lines": 3, "test_set_name": "t"})
    back = mat_io.load_processed_qsos(path)
    for n in mat_io.RESULT_VARIABLES:
        assert np.array_equal(back[n], res[n], equal_nan=True), n
    from scipy.io import loadmat
    assert loadmat(path)["min_z_dlas"].shape == (Q, 1)    # column vectors, as the script saves them


def test_v73_without_h5py_fails_loudly(tmp_path):
    try:
        import h5py  # noqa: F401
        pytest.skip("h5py is installed")
    except ImportError:
        pass
    p = tmp_path / "v73.mat"
    p.write_bytes(b"MATLAB 7.3 MAT-file".ljust(512, b" ") + b"\x89HDF\r\n\x1a\n" + b"\0" * 64)
    with pytest.raises(ImportError, match="h5py"):
        mat_io.load_dla_samples(str(p))
