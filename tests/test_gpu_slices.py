"""-m gpu: the exact-slice int8 tcgen05 contraction (csrc/slices_path.cuh), the default path for three lines
and the as-written gather (log_mvnpdf_low_rank.m:22, :26-28 off the FP64 pipe).

  * it IS the path that runs (dla_timing.contraction == 1) and it meets the north_star gates against the C oracle
  * it agrees with the FP64 DMMA kernels (DLA_B200_CONTRACT=dmma) far inside those gates
  * spectra shorter than a K step / a segment / a 16-pixel chunk (padding chunks, segments with no live pixel)
  * a quasar with huge dynamic range of the noise variance (per-pixel power-of-two scales, column scales)
  * NaN inputs poison exactly what they poison in fp64
  * chunked and overlapped schedules give bit-identical results
  * any rank 1 <= k <= 64 (column-unit layouts), BASELINE config 5 (k = 40)
  * batches the slice path must not take (masked gather, num_lines != 3) fall back to the DMMA kernels
All through the C ABI.
"""
import copy

import numpy as np
import pytest

from gp_dla_detection_b200.process_qsos import DlaContext, PackedBatch
from tests.helpers import ATOL_LOG_POST, RTOL_SAMPLE_LL, assert_parity, oracle_batch, prepared_config

pytestmark = pytest.mark.gpu


def _run(prepared, samples, k, S, lp_dla, lp_no, num_lines=3, workspace_bytes=0):
    with DlaContext(k=k, num_lines=num_lines, num_samples=S, workspace_bytes=workspace_bytes) as ctx:
        ctx.set_samples(samples.offset_samples, samples.nhi_samples)
        out = ctx.eval_batch(PackedBatch(prepared, k), lp_dla, lp_no)
        out["_timing"] = ctx.timing()
    return out


def test_slice_path_runs_and_meets_the_gates(oracle, monkeypatch):
    Q, S = 6, 1500   # 12 sample blocks of 128, the last one partial
    params, model, samples, sp, prepared = prepared_config("config1", Q, S)
    got = _run(prepared, samples, 20, S, sp.log_priors_dla, sp.log_priors_no_dla)
    assert got["_timing"]["contraction"] == 1, "the int8 slice contraction is not the path that ran"
    assert got["_timing"]["ms_chol"] > 0
    ref = oracle_batch(oracle, prepared, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    worst = assert_parity(got, ref)
    monkeypatch.setenv("DLA_B200_CONTRACT", "dmma")
    dmma = _run(prepared, samples, 20, S, sp.log_priors_dla, sp.log_priors_no_dla)
    assert dmma["_timing"]["contraction"] == 0
    a, b = got["sample_log_likelihoods_dla"], dmma["sample_log_likelihoods_dla"]
    both = float(np.max(np.abs(a - b)))
    print(f"slices vs oracle worst rel {worst:.2e}; slices vs DMMA worst abs {both:.2e}")
    assert both <= 2e-9     # the two contractions differ by rounding only (values are O(1e2..1e3))
    assert np.array_equal(got["map_ind"], dmma["map_ind"])


@pytest.mark.parametrize("n_keep", [1, 7, 15, 16, 17, 33, 64, 65, 129])
def test_short_spectra(oracle, n_keep):
    """Spectra of 1..129 kept pixels: fewer 16-pixel chunks than segments, chunks wholly beyond n, one K step."""
    S = 300
    params, model, samples, sp, prepared = prepared_config("config1", 2, S)
    short = []
    for p in prepared:
        p = copy.deepcopy(p)
        # keep the first n_keep pixels: every per-kept-pixel array and the padded grid shrink together
        p.flux, p.noise_variance, p.mu, p.omega2 = (x[:n_keep].copy() for x in (p.flux, p.noise_variance, p.mu, p.omega2))
        p.M = np.asfortranarray(np.asarray(p.M)[:n_keep])
        p.profile_index = p.profile_index[:n_keep].copy()
        p.padded_wavelengths = p.padded_wavelengths[:n_keep + 6].copy()
        p.n, p.n_full = n_keep, n_keep
        short.append(p)
    got = _run(short, samples, 20, S, sp.log_priors_dla, sp.log_priors_no_dla)
    assert got["_timing"]["contraction"] == 1
    ref = oracle_batch(oracle, short, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    assert_parity(got, ref)


def test_wide_dynamic_range_of_the_noise(oracle):
    """Noise variance from 1e-7 to 1e+3 inside ONE spectrum, |flux| up to 30: the per-pixel scales c_w, c_u span
    33 binary orders and the column scales follow the few low-noise pixels."""
    S = 400
    params, model, samples, sp, prepared = prepared_config("config1", 3, S)
    rng = np.random.default_rng(7)
    hard = []
    for p in prepared:
        p = copy.deepcopy(p)
        p.noise_variance = 10.0 ** rng.uniform(-7, 3, p.n)
        p.flux = p.flux * np.where(rng.random(p.n) < 0.02, 30.0, 1.0)
        hard.append(p)
    got = _run(hard, samples, 20, S, sp.log_priors_dla, sp.log_priors_no_dla)
    assert got["_timing"]["contraction"] == 1
    ref = oracle_batch(oracle, hard, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    worst = assert_parity(got, ref)
    print(f"noise variance 1e-7..1e3: worst per-sample rel err {worst:.2e}")


def test_nonfinite_inputs_poison_like_fp64(oracle):
    """A NaN flux pixel or a NaN entry of M makes every likelihood of THAT quasar NaN, and only of that quasar
    (the slices of a NaN are garbage integers: the NaN travels through r'D^-1 r and the column scales instead)."""
    S = 256
    params, model, samples, sp, prepared = prepared_config("config1", 4, S)
    bad = [copy.deepcopy(p) for p in prepared]
    bad[0].flux[100] = np.nan
    M1 = np.array(bad[1].M)
    M1[200, 3] = np.nan
    bad[1].M = np.asfortranarray(M1)
    bad[2].noise_variance[50] = np.nan
    got = _run(bad, samples, 20, S, sp.log_priors_dla, sp.log_priors_no_dla)
    assert got["_timing"]["contraction"] == 1
    ref = oracle_batch(oracle, bad, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    ll = got["sample_log_likelihoods_dla"]
    assert np.all(np.isnan(ll[:3]))
    assert np.all(np.isnan(got["log_likelihoods_no_dla"][:3]))
    assert np.all(np.isnan(ref["sample_log_likelihoods_dla"][:3]))
    rel = np.abs(ll[3] - ref["sample_log_likelihoods_dla"][3]) / np.abs(ref["sample_log_likelihoods_dla"][3])
    assert rel.max() <= RTOL_SAMPLE_LL


def test_schedules_are_bitwise_identical(monkeypatch):
    """One chunk, many chunks (small workspace) and the overlapped two-stream schedule (DLA_B200_OVERLAP=1: contraction
    of chunk i under the Voigt kernel of chunk i + 1, alternating plane sets) must agree bit for bit."""
    Q, S = 14, 700
    params, model, samples, sp, prepared = prepared_config("config3", Q, S)   # ragged lengths
    one = _run(prepared, samples, 20, S, sp.log_priors_dla, sp.log_priors_no_dla)
    assert one["_timing"]["contraction"] == 1
    many = _run(prepared, samples, 20, S, sp.log_priors_dla, sp.log_priors_no_dla, workspace_bytes=64 << 20)
    assert many["_timing"]["chunks"] > one["_timing"]["chunks"]
    monkeypatch.setenv("DLA_B200_OVERLAP", "1")
    ovl = _run(prepared, samples, 20, S, sp.log_priors_dla, sp.log_priors_no_dla)
    assert ovl["_timing"]["chunks"] >= 2
    monkeypatch.setenv("DLA_B200_K2S_STAGES", "3")
    monkeypatch.setenv("DLA_B200_OVERLAP", "0")
    st3 = _run(prepared, samples, 20, S, sp.log_priors_dla, sp.log_priors_no_dla)
    for other in (many, ovl, st3):
        for key in ("sample_log_likelihoods_dla", "log_likelihoods_no_dla", "log_posteriors_dla", "map_ind"):
            assert np.array_equal(one[key], other[key], equal_nan=True), key


@pytest.mark.parametrize("variant", ["0", "2"])
def test_voigt_kernel_builds(oracle, monkeypatch, variant):
    """DLA_B200_K1Q: the other builds of the slice-emitting Voigt kernel (default 1: four CTAs per SM, records in shared
    memory; 0: three CTAs per SM; 2: records from L1)."""
    monkeypatch.setenv("DLA_B200_K1Q", variant)
    params, model, samples, sp, prepared = prepared_config("config3", 4, 300)
    got = _run(prepared, samples, 20, 300, sp.log_priors_dla, sp.log_priors_no_dla)
    assert got["_timing"]["contraction"] == 1
    ref = oracle_batch(oracle, prepared, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    assert_parity(got, ref)


def test_fallbacks_to_the_dmma_kernels(oracle, monkeypatch):
    """Masked gather and num_lines != 3 are not slice-path batches: the FP64 kernels run them; so does any batch
    under DLA_B200_CONTRACT=dmma (k = 40: column-block DMMA contraction + shared-memory Cholesky)."""
    S = 200
    params, model, samples, sp, prepared = prepared_config("config3", 3, S, "intended")
    got = _run(prepared, samples, 20, S, sp.log_priors_dla, sp.log_priors_no_dla)
    assert got["_timing"]["contraction"] == 0
    params, model, samples, sp, prepared = prepared_config("config1", 2, S)
    got = _run(prepared, samples, 20, S, sp.log_priors_dla, sp.log_priors_no_dla, num_lines=5)
    assert got["_timing"]["contraction"] == 0
    params, model, samples, sp, prepared = prepared_config("config5", 2, S)
    monkeypatch.setenv("DLA_B200_CONTRACT", "dmma")
    got = _run(prepared, samples, 40, S, sp.log_priors_dla, sp.log_priors_no_dla)
    assert got["_timing"]["contraction"] == 0
    ref = oracle_batch(oracle, prepared, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    assert_parity(got, ref)


def test_rank40_config5_takes_the_slice_path(oracle, monkeypatch):
    """BASELINE config 5 (k = 40, 2433 px): 13 + 1 column units of 64, rows of G 896 wide, generic bordered Cholesky."""
    S = 300
    params, model, samples, sp, prepared = prepared_config("config5", 3, S)
    got = _run(prepared, samples, 40, S, sp.log_priors_dla, sp.log_priors_no_dla)
    assert got["_timing"]["contraction"] == 1
    ref = oracle_batch(oracle, prepared, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    worst = assert_parity(got, ref)
    monkeypatch.setenv("DLA_B200_CONTRACT", "dmma")
    dmma = _run(prepared, samples, 40, S, sp.log_priors_dla, sp.log_priors_no_dla)
    both = float(np.max(np.abs(got["sample_log_likelihoods_dla"] - dmma["sample_log_likelihoods_dla"])))
    print(f"k = 40: slices vs oracle worst rel {worst:.2e}; slices vs DMMA worst abs {both:.2e}")
    assert both <= 5e-9
    assert np.array_equal(got["map_ind"], dmma["map_ind"])


@pytest.mark.parametrize("k", [1, 7, 21, 32, 33, 45, 64])
def test_any_rank_on_the_slice_path(oracle, k):
    """Column-unit layouts of every kind: k <= 20 fixed layout; k = 21 (231 = 3 x 64 + 39 -> four 64-units | 32),
    32 (528 = 8 x 64 + 16 -> a 32-unit | 32), 33 (561 -> 49 left: ninth 64-unit | 64), 45 (1035 -> 11 left: 32-unit | 64),
    64 (2080 = 32 x 64 + 32 -> 32-unit | 64)."""
    S = 200
    params, model, samples, sp, prepared = prepared_config("config3", 3, S)
    rng = np.random.default_rng(100 + k)
    mine = []
    for p in prepared:
        p = copy.deepcopy(p)
        n = p.flux.shape[0]
        p.M = np.asfortranarray(rng.normal(size=(n, k)) * (0.4 / np.sqrt(k)) * (1 + np.arange(k)) ** -0.3)
        mine.append(p)
    got = _run(mine, samples, k, S, sp.log_priors_dla, sp.log_priors_no_dla)
    assert got["_timing"]["contraction"] == 1
    ref = oracle_batch(oracle, mine, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    assert_parity(got, ref)


def test_spectra_longer_than_the_int32_accumulator_bound_take_the_fp64_kernels(oracle):
    """The int32 slice accumulators are exact up to 9 362 pixels; the planner sends quasars of more than 8 192 kept pixels
    to the FP64 kernels (here: the model on a 0.03 A grid, 10 134 pixels, twice the pipeline's finest)."""
    from gp_dla_detection_b200 import synthetic
    from gp_dla_detection_b200.process_qsos import prepare_quasar
    from gp_dla_detection_b200.set_parameters import Parameters
    S, k = 100, 20
    params = Parameters(k=k, dlambda=0.03, num_dla_samples=S)
    model = synthetic.make_model(k, 0.03, 20260077, params)
    samples = synthetic.make_samples(S, params)
    sp = synthetic.make_spectra(2, model, 20260078, False, params)
    prepared = [prepare_quasar(w, f, v, m, z, model, params)
                for w, f, v, m, z in zip(sp.all_wavelengths, sp.all_flux, sp.all_noise_variance, sp.all_pixel_mask, sp.z_qsos)]
    assert min(p.n for p in prepared) > 8192
    got = _run(prepared, samples, k, S, sp.log_priors_dla, sp.log_priors_no_dla)
    assert got["_timing"]["contraction"] == 0
    ref = oracle_batch(oracle, prepared, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    assert_parity(got, ref)


def test_a_batch_whose_chunks_take_different_kernel_families(oracle):
    """Masked quasars with the intended gather (config 3: gather != identity -> FP64 kernels) interleaved with unmasked
    ones (config 1 -> int8 slice path), one quasar per chunk, inside ONE call: every quasar meets the gates, and equals
    what it gives when the whole batch is one (FP64) chunk to rounding."""
    S = 256
    _, _, samples, sp3, prep3 = prepared_config("config3", 4, S, "intended")
    _, _, samples1, sp1, prep1 = prepared_config("config1", 4, S, "intended")
    assert np.array_equal(samples.nhi_samples, samples1.nhi_samples)
    prepared = [p for pair in zip(prep3, prep1) for p in pair]
    lpd = np.array([v for pair in zip(sp3.log_priors_dla, sp1.log_priors_dla) for v in pair])
    lpn = np.array([v for pair in zip(sp3.log_priors_no_dla, sp1.log_priors_no_dla) for v in pair])
    ident = [bool(np.array_equal(p.profile_index, np.arange(p.n))) for p in prepared]
    assert ident == [False, True] * 4
    split = _run(prepared, samples, 20, S, lpd, lpn, workspace_bytes=16 * 320 * 1232)
    assert split["_timing"]["chunks"] == 8
    ref = oracle_batch(oracle, prepared, samples, lpd, lpn)
    assert_parity(split, ref)
    whole = _run(prepared, samples, 20, S, lpd, lpn)
    assert whole["_timing"]["chunks"] == 1 and whole["_timing"]["contraction"] == 0
    assert np.max(np.abs(split["sample_log_likelihoods_dla"] - whole["sample_log_likelihoods_dla"])) <= 2e-9
    assert np.array_equal(split["map_ind"], whole["map_ind"])
