"""ASCII catalogue of a DLA search: the three files ``generate_ascii_catalog.m`` writes.

Formats follow the reference's ``fprintf`` calls conversion by conversion
(generate_ascii_catalog.m:15, :29, :60, :63-74, :82); MATLAB prints non-finite values as
``Inf`` / ``-Inf`` / ``NaN``, which Python's ``%`` does not, hence ``_m``.  The MAP sample
(:76-80) comes from the device (``map_ind`` = first index of the NaN-ignoring maximum of
``sample_log_likelihoods_dla``) or, when only the likelihood matrix is given, is recomputed
here with the same tie rule.
"""
from __future__ import annotations

import math
import re
from typing import Iterable, Sequence

import numpy as np


def _m(fmt: str, x: float) -> str:
    """One MATLAB numeric conversion (``%w.pf`` / ``%w.pe``) including its non-finite spellings."""
    x = float(x)
    if math.isfinite(x):
        return fmt % x
    width = re.match(r"%[-+ 0]*(\d*)", fmt).group(1)
    text = "NaN" if math.isnan(x) else ("Inf" if x > 0 else "-Inf")
    if "+" in fmt and x > 0:
        text = "+" + text
    return text.rjust(int(width)) if width else text


def _e3(x: float) -> str:
    """sprintf('%0.5e', x) with the exponent widened to three digits (generate_ascii_catalog.m:71-74)."""
    return re.sub(r"e([+-])(\d\d)$", r"e\g<1>0\2", _m("%0.5e", x))


def format_dla_sample(offset: float, log_nhi: float) -> str:
    return "%s %s\n" % (_m("%06f", offset), _m("%09f", log_nhi))                    # :15


def format_spectrum(thing_id: int, sdss_name: str, plate: int, mjd: int, fiber_id: int, ra: float, dec: float,
                    z_qso: float, snr: float, filter_flags: int) -> str:
    bits = [(int(filter_flags) >> b) & 1 for b in range(4)]                          # bitget(flags, 1..4)
    return ("%09d %-18s %04d %05d %04d %s %s %s %s %d%d%d%d\n"                        # :29
            % (int(thing_id), str(sdss_name).rstrip(), int(plate), int(mjd), int(fiber_id), _m("%011.7f", ra),
               _m("%+011.7f", dec), _m("%06.4f", z_qso), _m("%08.4f", snr), *bits))


def format_result(thing_id: int, min_z_dla: float, max_z_dla: float, log_prior_no_dla: float, log_prior_dla: float,
                  log_likelihood_no_dla: float, log_likelihood_dla: float, p_no_dla: float, p_dla: float,
                  map_z_dla: float, map_log_nhi: float) -> str:
    # :60 passes ONE argument to '%09i %-18s ': MATLAB stops at the first conversion without data,
    # so only the id and the blank after it are written
    head = "%09d " % int(thing_id)
    body = "%s %s %s %s %s %s %s %s " % (                                             # :63-74
        _m("%06.4f", min_z_dla), _m("%06.4f", max_z_dla), _m("%8.5f", log_prior_no_dla), _m("%8.5f", log_prior_dla),
        _m("%12.5e", log_likelihood_no_dla), _m("%12.5e", log_likelihood_dla), _e3(p_no_dla), _e3(p_dla))
    tail = "%s %s\n" % (_m("%06.4f", map_z_dla), _m("%07.4f", map_log_nhi))            # :82
    return head + body + tail


def map_sample(sample_log_likelihoods_dla: np.ndarray) -> np.ndarray:
    """[~, map_ind] = nanmax(row): first index of the maximum, NaN ignored; -1 for an all-NaN row."""
    ll = np.asarray(sample_log_likelihoods_dla, dtype=np.float64)
    out = np.full(ll.shape[0], -1, dtype=np.int64)
    for i, row in enumerate(ll):
        ok = ~np.isnan(row)
        if ok.any():
            out[i] = int(np.flatnonzero(ok & (row == row[ok].max()))[0])
    return out


def write_dla_samples(path: str, offset_samples: Iterable[float], log_nhi_samples: Iterable[float]) -> None:
    with open(path, "w") as f:
        for o, l in zip(offset_samples, log_nhi_samples):
            f.write(format_dla_sample(o, l))


def write_spectra(path: str, catalog: dict) -> None:
    """catalog: thing_ids, sdss_names, plates, mjds, fiber_ids, ras, decs, z_qsos, snrs, filter_flags."""
    with open(path, "w") as f:
        for i in range(len(catalog["z_qsos"])):
            f.write(format_spectrum(catalog["thing_ids"][i], catalog["sdss_names"][i], catalog["plates"][i],
                                    catalog["mjds"][i], catalog["fiber_ids"][i], catalog["ras"][i], catalog["decs"][i],
                                    catalog["z_qsos"][i], catalog["snrs"][i], catalog["filter_flags"][i]))


def write_results(path: str, thing_ids: Sequence[int], results: dict, offset_samples, log_nhi_samples) -> None:
    """results: the dict process_qsos() returns (variable names of process_qsos.m:235-243);
    thing_ids: catalog.thing_ids(find(test_ind))."""
    off = np.asarray(offset_samples, dtype=np.float64)
    lnhi = np.asarray(log_nhi_samples, dtype=np.float64)
    mi = results.get("map_ind")
    if mi is None:
        mi = map_sample(results["sample_log_likelihoods_dla"])
    with open(path, "w") as f:
        for i, tid in enumerate(thing_ids):
            j = int(mi[i])
            if j >= 0:
                mz = results["min_z_dlas"][i] + (results["max_z_dlas"][i] - results["min_z_dlas"][i]) * off[j]   # :78-79
                mn = lnhi[j]
            else:
                mz = mn = float("nan")
            f.write(format_result(tid, results["min_z_dlas"][i], results["max_z_dlas"][i],
                                  results["log_priors_no_dla"][i], results["log_priors_dla"][i],
                                  results["log_likelihoods_no_dla"][i], results["log_likelihoods_dla"][i],
                                  results["model_posteriors"][i][0], results["model_posteriors"][i][1], mz, mn))
