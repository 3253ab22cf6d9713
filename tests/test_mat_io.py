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


@pytest.mark.parametrize("v73", [True, False])
def test_processed_qsos_round_trip(tmp_path, v73):
    Q, S = 4, 9
    rng = np.random.default_rng(1)
    res = {n: rng.standard_normal(Q) for n in mat_io.RESULT_VARIABLES}
    res["sample_log_likelihoods_dla"] = rng.standard_normal((Q, S))
    res["sample_log_likelihoods_dla"][1, 3] = np.nan
    res["model_posteriors"] = rng.random((Q, 2))
    res["log_priors_dla"][0] = -np.inf
    path = str(tmp_path / "processed_qsos_t.mat")
    extra = {"num_lines": 3, "test_set_name": "dr12q", "test_ind": np.array([True, False, True, True]),
             "max_z_cut": 0.0100069}
    mat_io.save_processed_qsos(path, res, extra=extra, v73=v73)
    assert mat_io._is_hdf5(path) == v73
    back = mat_io.load_processed_qsos(path)
    for n in mat_io.RESULT_VARIABLES:
        assert np.array_equal(back[n], res[n], equal_nan=True), n
    if not v73:
        from scipy.io import loadmat
        assert loadmat(path)["min_z_dlas"].shape == (Q, 1)    # column vectors, as the script saves them
        return
    from gp_dla_detection_b200 import hdf5_min
    with hdf5_min.File(path) as f:                            # the bookkeeping variables of process_qsos.m:235-238
        assert f.read("test_set_name") == "dr12q" and f.attrs("test_set_name") == {"MATLAB_class": b"char", "MATLAB_int_decode": 2}
        assert np.array_equal(f.read("test_ind").ravel(), extra["test_ind"]) and f.attrs("test_ind")["MATLAB_class"] == b"logical"
        assert f.read("num_lines")[0, 0] == 3.0 and f.attrs("num_lines") == {"MATLAB_class": b"double"}
        assert f.read("min_z_dlas").shape == (Q, 1) and f.read("sample_log_likelihoods_dla").shape == (Q, S)


def test_v73_default_and_v5_limit(tmp_path, monkeypatch):
    """process_qsos.m:249 saves -v7.3; that is the default here, and a result the v5 format cannot hold is refused."""
    import inspect
    assert inspect.signature(mat_io.save_processed_qsos).parameters["v73"].default is True
    monkeypatch.setattr(mat_io, "V5_LIMIT_BYTES", 64)
    with pytest.raises(ValueError, match="v73=True"):
        mat_io.save_processed_qsos(str(tmp_path / "x.mat"), {"sample_log_likelihoods_dla": np.zeros((3, 9))}, v73=False)


def test_v73_cell_arrays_round_trip(tmp_path):
    """preloaded_qsos.mat as v7.3: ragged cell arrays are datasets of object references into #refs#."""
    from gp_dla_detection_b200 import hdf5_min
    params, model, samples, sp = synthetic.make_config("config3", num_quasars=5, num_samples=8)
    path = str(tmp_path / "preloaded_qsos.mat")
    hdf5_min.write(path, {"all_wavelengths": list(sp.all_wavelengths), "all_flux": list(sp.all_flux),
                          "all_noise_variance": list(sp.all_noise_variance),
                          "all_pixel_mask": [np.asarray(m, dtype=bool) for m in sp.all_pixel_mask]})
    p = mat_io.load_preloaded_qsos(path)
    for q in range(5):
        assert np.array_equal(p["all_wavelengths"][q], sp.all_wavelengths[q])
        assert np.array_equal(p["all_noise_variance"][q], sp.all_noise_variance[q])
        assert np.array_equal(p["all_pixel_mask"][q], sp.all_pixel_mask[q]) and p["all_pixel_mask"][q].dtype == bool
    mpath = str(tmp_path / "learned_qso_model_x.mat")
    hdf5_min.write(mpath, {"rest_wavelengths": model.rest_wavelengths, "mu": model.mu, "M": model.M, "log_omega": model.log_omega,
                           "log_c_0": model.log_c_0, "log_tau_0": model.log_tau_0, "log_beta": model.log_beta})
    m2 = mat_io.load_learned_model(mpath)
    assert np.array_equal(m2.M, model.M) and m2.log_beta == model.log_beta and np.array_equal(m2.mu, model.mu)


def test_v73_chunked_dataset_is_refused_with_a_clear_message(tmp_path):
    """A MATLAB-written file with a chunked (compressed) variable: outside hdf5_min's subset; say what is needed."""
    from gp_dla_detection_b200 import hdf5_min
    path = str(tmp_path / "c.mat")
    hdf5_min.write(path, {"x": np.arange(6.0).reshape(2, 3)})
    raw = bytearray(open(path, "rb").read())
    at = raw.index(bytes([8, 0, 24, 0, 0, 0, 0, 0, 3, 1]))       # the data-layout message: type 8, size 24, version 3, class 1
    raw[at + 9] = 2                                              # class 2 = chunked
    open(path, "wb").write(bytes(raw))
    try:
        import h5py  # noqa: F401
        pytest.skip("h5py is installed")
    except ImportError:
        pass
    with pytest.raises(hdf5_min.Hdf5FormatError, match="h5py"):
        mat_io._load(path, ("x",))

