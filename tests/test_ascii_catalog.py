"""generate_ascii_catalog.m formats (:15, :29, :60-82): known-answer lines worked out by hand from
the fprintf conversions, MATLAB's non-finite spellings, and the first-index nanmax of the MAP sample."""
import numpy as np

from gp_dla_detection_b200 import generate_ascii_catalog as G


def test_sample_and_spectrum_lines():
    assert G.format_dla_sample(0.5, 20.25) == "0.500000 20.250000\n"
    assert G.format_dla_sample(0.000123, 9.5) == "0.000123 09.500000\n"       # %06f never pads here, %09f does
    line = G.format_spectrum(87654321, "003000.12+001122.3  ", 266, 51630, 7, 7.50050, -1.189833, 2.4513, 12.3456, 0b0101)
    assert line == "087654321 003000.12+001122.3 0266 51630 0007 007.5005000 -01.1898330 2.4513 012.3456 1010\n"
    line = G.format_spectrum(1, "x", 12345, 5, 999, 359.9, 12.0, 10.5, 7.0, 8)
    assert line == "000000001 x                  12345 00005 0999 359.9000000 +12.0000000 10.5000 007.0000 0001\n"


def test_result_line_including_matlab_quirks():
    # one argument for '%09i %-18s ' -> only the id and a blank (generate_ascii_catalog.m:60)
    line = G.format_result(42, 2.01234, 3.5, np.log(0.9), np.log(0.1), -1234.5678, -1200.25, 1.5e-15, 1.0, 2.98766, 20.5)
    assert line == ("000000042 2.0123 3.5000 -0.10536 -2.30259 -1.23457e+03 -1.20025e+03 "
                    "1.50000e-015 1.00000e+000 2.9877 20.5000\n")
    # -Inf prior (process_qsos.m:129-132) and NaN results (:74-82 initialise with NaN)
    line = G.format_result(7, 2.0, 2.5, 0.0, -np.inf, np.nan, -5.0, 1.0, 0.0, np.nan, np.nan)
    assert line == ("000000007 2.0000 2.5000  0.00000     -Inf          NaN -5.00000e+00 "
                    "1.00000e+000 0.00000e+000    NaN     NaN\n")


def test_map_sample_is_first_index_nanmax(tmp_path):
    ll = np.array([[1.0, np.nan, 3.0, 3.0], [np.nan, np.nan, np.nan, np.nan], [-np.inf, -2.0, np.nan, -2.0]])
    assert list(G.map_sample(ll)) == [2, -1, 1]
    res = {"sample_log_likelihoods_dla": ll, "min_z_dlas": [2.0, 2.0, 2.0], "max_z_dlas": [3.0, 3.0, 3.0],
           "log_priors_no_dla": [np.log(0.9)] * 3, "log_priors_dla": [np.log(0.1)] * 3,
           "log_likelihoods_no_dla": [-10.0] * 3, "log_likelihoods_dla": [-9.0] * 3,
           "model_posteriors": [[0.75, 0.25]] * 3}
    off, lnhi = [0.0, 0.25, 0.5, 0.75], [20.0, 20.5, 21.0, 21.5]
    p = tmp_path / "r.dat"
    G.write_results(str(p), [1, 2, 3], res, off, lnhi)
    rows = p.read_text().splitlines()
    assert rows[0].endswith("7.50000e-001 2.50000e-001 2.5000 21.0000")
    assert rows[1].endswith("   NaN     NaN")
    assert rows[2].endswith("2.2500 20.5000")
    G.write_dla_samples(str(tmp_path / "s.dat"), off, lnhi)
    assert (tmp_path / "s.dat").read_text().splitlines()[1] == "0.250000 20.500000"
