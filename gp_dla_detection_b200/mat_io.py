"""On-disk products either side of the hot path, by the reference's variable names.

  learned_qso_model_<set>.mat  learn_qso_model.m:113-123   -> load_learned_model
  dla_samples.mat              generate_dla_samples.m:59-63 -> load_dla_samples
  preloaded_qsos.mat           preload_qsos.m:73-79         -> load_preloaded_qsos
  processed_qsos_<set>.mat     process_qsos.m:235-249       -> save_processed_qsos / load_processed_qsos

The reference writes MATLAB v7.3 files (= HDF5; process_qsos.m:249, because sample_log_likelihoods_dla is
[num_quasars x 10000] doubles: 13 GB for DR12Q).  There is no h5py in this image, so the subset of HDF5 these
files need is written and read by hdf5_min.py (numpy + struct: contiguous little-endian datasets, MATLAB's
transposed layout, MATLAB_class attributes, cell arrays as object references -- the conventions
CDDF_analysis/calc_cddf.py:59-124 relies on).  save_processed_qsos writes v7.3 by default, as the reference does.
Reading MATLAB-WRITTEN v7.3 files whose variables are chunked / compressed (MATLAB's default for large arrays):
needs h5py, unavailable on this box -- hdf5_min says so when it meets one, and h5py is used when importable.
The v5 / v7 format (scipy.io; plain ``load`` in MATLAB and Octave) remains for small files and cannot hold a
variable of 2 GiB or more: asking for it with such a result is an error, not a silent truncation.
"""
from __future__ import annotations

import numpy as np

from .process_qsos import DLASamples, LearnedModel

RESULT_VARIABLES = ("min_z_dlas", "max_z_dlas", "log_priors_no_dla", "log_priors_dla", "log_likelihoods_no_dla",
                    "sample_log_likelihoods_dla", "log_likelihoods_dla", "log_posteriors_no_dla", "log_posteriors_dla",
                    "model_posteriors", "p_no_dlas", "p_dlas")  # process_qsos.m:235-243


def _is_hdf5(path: str) -> bool:
    with open(path, "rb") as f:
        head = f.read(512 + 8)
    return head[:8] == b"\x89HDF\r\n\x1a\n" or head[512:520] == b"\x89HDF\r\n\x1a\n"


def _h5py():
    try:
        import h5py  # noqa: F401
        return h5py
    except ImportError:
        return None


V5_LIMIT_BYTES = 2 ** 31 - 8   # a MATLAB v5 / v7 data element carries a 32-bit byte count


def _load(path: str, names) -> dict:
    """name -> ndarray (numeric) or list of 1-D ndarrays (cell array of vectors), MATLAB orientation."""
    if _is_hdf5(path):
        h5py = _h5py()
        out = {}
        if h5py is None:
            from . import hdf5_min
            with hdf5_min.File(path) as f:
                for n in names:
                    if n not in f.members:
                        raise KeyError(f"{path}: no variable {n!r}")
                    out[n] = f.read(n)
            return out
        with h5py.File(path, "r") as f:
            for n in names:
                ds = f[n]
                if ds.dtype == object:   # cell array: references to the per-quasar vectors
                    out[n] = [np.asarray(f[r]).ravel() for r in np.asarray(ds).ravel()]
                else:
                    out[n] = np.asarray(ds).T  # HDF5 stores MATLAB arrays transposed
        return out
    from scipy.io import loadmat
    raw = loadmat(path, variable_names=list(names))
    out = {}
    for n in names:
        v = raw[n]
        if v.dtype == object:
            out[n] = [np.asarray(x).ravel() for x in v.ravel()]
        else:
            out[n] = np.asarray(v)
    return out


def load_learned_model(path: str) -> LearnedModel:
    v = _load(path, ("rest_wavelengths", "mu", "M", "log_omega", "log_c_0", "log_tau_0", "log_beta"))  # process_qsos.m:30-35
    rest = np.asarray(v["rest_wavelengths"], dtype=np.float64).ravel()
    M = np.asarray(v["M"], dtype=np.float64)
    if M.shape[0] != rest.shape[0]:
        M = M.T
    return LearnedModel(rest, np.asarray(v["mu"], dtype=np.float64).ravel(), np.asfortranarray(M),
                        np.asarray(v["log_omega"], dtype=np.float64).ravel(), float(np.ravel(v["log_c_0"])[0]),
                        float(np.ravel(v["log_tau_0"])[0]), float(np.ravel(v["log_beta"])[0]))


def load_dla_samples(path: str) -> DLASamples:
    v = _load(path, ("offset_samples", "log_nhi_samples", "nhi_samples"))                # process_qsos.m:37-40
    return DLASamples(*(np.asarray(v[n], dtype=np.float64).ravel() for n in ("offset_samples", "log_nhi_samples", "nhi_samples")))


def load_preloaded_qsos(path: str) -> dict:
    """all_wavelengths / all_flux / all_noise_variance / all_pixel_mask as lists of 1-D arrays (process_qsos.m:42-50)."""
    v = _load(path, ("all_wavelengths", "all_flux", "all_noise_variance", "all_pixel_mask"))
    v["all_pixel_mask"] = [np.asarray(m).astype(bool) for m in v["all_pixel_mask"]]
    return v


def save_processed_qsos(path: str, results: dict, extra: dict | None = None, v73: bool = True) -> None:
    """processed_qsos_<set>.mat with the variables of process_qsos.m:235-243 (column vectors, [Q x S]
    sample matrix, [Q x 2] posteriors).  ``extra``: the bookkeeping variables of :235-238 (release names,
    test_ind, ...).  v73=True (default, as process_qsos.m:249): HDF5 with MATLAB's transposed layout and
    MATLAB_class attributes, streamed by hdf5_min (no h5py needed).  v73=False: MATLAB v5 through scipy.io;
    raises when a variable does not fit that format."""
    vars_ = {}
    for n in RESULT_VARIABLES:
        if results.get(n) is None:
            continue
        a = np.asarray(results[n], dtype=np.float64)
        vars_[n] = a.reshape(-1, 1) if a.ndim == 1 else a
    for n, a in (extra or {}).items():
        vars_[n] = a
    if v73:
        from . import hdf5_min
        hdf5_min.write(path, vars_)
        return
    for n, a in vars_.items():
        nbytes = np.asarray(a).nbytes if not isinstance(a, str) else len(a)
        if nbytes > V5_LIMIT_BYTES:
            raise ValueError(f"{n} is {nbytes} bytes: a MATLAB v5 file cannot hold it; save with v73=True "
                             f"(the reference's own format, process_qsos.m:249)")
    from scipy.io import savemat
    savemat(path, vars_, do_compression=False, oned_as="column")


def load_processed_qsos(path: str) -> dict:
    v = _load(path, RESULT_VARIABLES)
    for n in RESULT_VARIABLES:
        if n not in ("sample_log_likelihoods_dla", "model_posteriors"):
            v[n] = np.asarray(v[n]).ravel()
    return v
