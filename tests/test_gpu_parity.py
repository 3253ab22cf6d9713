"""-m gpu: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs."""
import numpy as np
import pytest

from gp_dla_detection_b200.process_qsos import DlaContext, PackedBatch, finish_outputs, process_qsos
from tests.helpers import assert_parity, oracle_batch, prepared_config

pytestmark = pytest.mark.gpu


def _run_gpu(prepared, samples, k, S, lp_dla, lp_no, num_lines=3, **kw):
    with DlaContext(k=k, num_lines=num_lines, num_samples=S, **kw) as ctx:
        ctx.set_samples(samples.offset_samples, samples.nhi_samples)
        out = ctx.eval_batch(PackedBatch(prepared, k), lp_dla, lp_no)
        out["_timing"] = ctx.timing()
    return out


def test_voigt_seam_matches_reference_voigt_c(oracle):
    """dla_voigt == the voigt MEX (voigt.c verbatim where built, else the C restatement)."""
    params, model, samples, sp, prepared = prepared_config("config1", 2, 8)
    ref_fn = oracle.voigt_ref if oracle.load_ref() is not None else oracle.voigt_c
    with DlaContext(k=20, num_samples=8) as ctx:
        for p in prepared:
            for frac, N in [(0.5, 0.0), (0.1, 1e20), (0.5, 10 ** 20.5), (0.9, 1e21), (0.3, 1e22), (0.7, 1e23)]:
                z = p.min_z_dla + (p.max_z_dla - p.min_z_dla) * frac
                got = ctx.voigt(p.padded_wavelengths, z, N, 3)
                ref = ref_fn(p.padded_wavelengths, z, N, 3)
                assert got.shape == ref.shape == (p.n_full,)
                assert np.max(np.abs(got - ref)) <= 1e-12
        # K2: N = 0 gives exactly 1.0 everywhere (sum of taps in tap order)
        assert np.all(ctx.voigt(prepared[0].padded_wavelengths, 3.0, 0.0, 3) == 1.0)
        # all 31 lines (the MEX default, voigt.c:16)
        p = prepared[0]
        z = p.min_z_dla + 0.4 * (p.max_z_dla - p.min_z_dla)
        assert np.max(np.abs(ctx.voigt(p.padded_wavelengths, z, 1e21, None) - ref_fn(p.padded_wavelengths, z, 1e21, 31))) <= 1e-12


@pytest.mark.parametrize("n,k", [(64, 1), (97, 5), (256, 20), (1217, 20), (1000, 13), (300, 21), (2433, 40), (777, 64)])
def test_log_mvnpdf_low_rank_seam(oracle, n, k):
    rng = np.random.default_rng(n * 100 + k)
    y, mu = rng.standard_normal(n), rng.standard_normal(n) * 0.1
    M = rng.standard_normal((n, k)) * 0.5
    d = np.exp(rng.standard_normal(n) * 0.7) * 0.05
    with DlaContext(k=20, num_samples=8) as ctx:
        got = ctx.log_mvnpdf_low_rank(y, mu, M, d)
    ref = oracle.log_mvnpdf_low_rank_c(y, mu, M, d)
    assert abs(got - ref) <= 1e-10 * abs(ref)
    if n <= 256:
        assert abs(got - oracle.log_mvnpdf_dense(y, mu, M, d)) <= 1e-9 * abs(ref)


def test_tiny_config_parity(oracle):
    params, model, samples, sp, prepared = prepared_config("tiny", 4, 64)
    got = _run_gpu(prepared, samples, 5, 64, sp.log_priors_dla, sp.log_priors_no_dla)
    ref = oracle_batch(oracle, prepared, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    assert_parity(got, ref)


def test_config1_subset_parity(oracle):
    """config 1 shape (n = 1217, k = 20) on a sample subset the oracle finishes in seconds."""
    params, model, samples, sp, prepared = prepared_config("config1", 3, 700)  # S not a multiple of 64
    got = _run_gpu(prepared, samples, 20, 700, sp.log_priors_dla, sp.log_priors_no_dla)
    ref = oracle_batch(oracle, prepared, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    worst = assert_parity(got, ref)
    print("config1 subset worst per-sample rel err", worst)


def test_config5_rank40_parity(oracle):
    """BASELINE config 5 shape (k = 40 null model on the 0.125 A grid, 2433 pixels): the column-block
    contraction + shared-memory Cholesky path, on a subset the oracle finishes in seconds."""
    params, model, samples, sp, prepared = prepared_config("config5", 2, 200)
    assert prepared[0].n == 2433 and prepared[0].M.shape[1] == 40
    got = _run_gpu(prepared, samples, 40, 200, sp.log_priors_dla, sp.log_priors_no_dla)
    ref = oracle_batch(oracle, prepared, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    worst = assert_parity(got, ref)
    print("config5 subset worst per-sample rel err", worst)


@pytest.mark.parametrize("mask_gather", ["reference", "intended"])
def test_ragged_config3_parity(oracle, mask_gather):
    params, model, samples, sp, prepared = prepared_config("config3", 6, 192, mask_gather)
    assert len({p.n for p in prepared}) > 1
    got = _run_gpu(prepared, samples, 20, 192, sp.log_priors_dla, sp.log_priors_no_dla)
    ref = oracle_batch(oracle, prepared, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    assert_parity(got, ref)


def test_chunking_and_staged_api_are_bitwise_identical():
    params, model, samples, sp, prepared = prepared_config("config3", 7, 256)
    k, S = 20, 256
    a = _run_gpu(prepared, samples, k, S, sp.log_priors_dla, sp.log_priors_no_dla)
    # a workspace that fits one quasar at a time -> 7 chunks
    b = _run_gpu(prepared, samples, k, S, sp.log_priors_dla, sp.log_priors_no_dla, workspace_bytes=16 * 320 * 1232)
    assert b["_timing"]["chunks"] > 1 and a["_timing"]["chunks"] == 1
    with DlaContext(k=k, num_samples=S, store_sample_ll=False) as ctx:
        ctx.set_samples(samples.offset_samples, samples.nhi_samples)
        ctx.stage(PackedBatch(prepared, k), sp.log_priors_dla, sp.log_priors_no_dla)
        ctx.run_staged()
        ctx.run_staged()
        c = ctx.fetch()
    for key in ("log_likelihoods_no_dla", "log_likelihoods_dla", "log_posteriors_dla", "model_posteriors", "map_ind"):
        assert np.array_equal(a[key], b[key]), key
        assert np.array_equal(a[key], c[key]), key
    assert np.array_equal(a["sample_log_likelihoods_dla"], b["sample_log_likelihoods_dla"])
    assert c["sample_log_likelihoods_dla"] is None


def test_edge_semantics(oracle):
    """-inf priors (process_qsos.m:129-132), N = 0 sample == no-DLA likelihood (K7),
    unguarded min_z > max_z (:160-165), tiny n."""
    params, model, samples, sp, prepared = prepared_config("config1", 2, 64)
    off = samples.offset_samples.copy()
    nhi = samples.nhi_samples.copy()
    nhi[5] = 0.0
    prepared[1].min_z_dla, prepared[1].max_z_dla = prepared[1].max_z_dla, prepared[1].min_z_dla
    lp_dla = np.array([-np.inf, np.log(0.3)])
    lp_no = np.array([0.0, np.log(0.7)])
    with DlaContext(k=20, num_samples=64) as ctx:
        ctx.set_samples(off, nhi)
        got = ctx.eval_batch(PackedBatch(prepared, 20), lp_dla, lp_no)
    assert got["sample_log_likelihoods_dla"][0, 5] == pytest.approx(got["log_likelihoods_no_dla"][0], rel=1e-13)
    assert got["log_posteriors_dla"][0] == -np.inf
    assert got["model_posteriors"][0, 0] == 1.0 and got["model_posteriors"][0, 1] == 0.0

    class S_:  # same samples object shape as DLASamples
        offset_samples, nhi_samples = off, nhi
    ref = oracle_batch(oracle, prepared, S_, lp_dla, lp_no)
    assert_parity(got, ref)


def test_full_size_properties():
    """BASELINE config-1 shape at full S = 10000: size-independent properties (K7, K10)."""
    params, model, samples, sp, prepared = prepared_config("config1", 4, 10000)
    S = 10000
    got = _run_gpu(prepared, samples, 20, S, sp.log_priors_dla, sp.log_priors_no_dla)
    ll = got["sample_log_likelihoods_dla"]
    assert np.all(np.isfinite(ll))
    # K10 (calc_cddf.py:98, :244-246): normalised sample likelihoods sum to 1
    norm = np.exp(ll - (got["log_likelihoods_dla"][:, None] + np.log(S))).sum(axis=1)
    assert np.allclose(norm, 1.0, rtol=1e-10)
    assert np.array_equal(got["map_ind"], np.argmax(ll, axis=1))
    assert np.allclose(got["model_posteriors"].sum(axis=1), 1.0, rtol=1e-14)
    # injected DLAs are found, clean sightlines are not
    p_dla = 1 - got["model_posteriors"][:, 0]
    print("p_dla", p_dla, "injected", sp.injected)


def test_process_qsos_drop_in(oracle):
    params, model, samples, sp, prepared = prepared_config("tiny", 4, 64)
    out = process_qsos(sp.all_wavelengths, sp.all_flux, sp.all_noise_variance, sp.all_pixel_mask, sp.z_qsos, model,
                       samples, sp.log_priors_dla, sp.log_priors_no_dla, params)
    for key in ("min_z_dlas", "max_z_dlas", "log_priors_no_dla", "log_priors_dla", "log_likelihoods_no_dla",
                "sample_log_likelihoods_dla", "log_likelihoods_dla", "log_posteriors_no_dla", "log_posteriors_dla",
                "model_posteriors", "p_no_dlas", "p_dlas"):  # process_qsos.m:235-243
        assert key in out
    assert np.allclose(out["p_dlas"], 1 - out["model_posteriors"][:, 0])
    ref = oracle_batch(oracle, prepared, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    assert_parity(out, ref)


def test_gpu_against_committed_goldens():
    """The CUDA path against tests/golden/golden_v1.npz (voigt.c verbatim profiles; numpy / C+voigt.c
    restatement log-likelihoods) -- nothing from oracle/ is executed here."""
    import os
    from tests.golden.make_golden import checksum
    G = np.load(os.path.join(os.path.dirname(__file__), "golden", "golden_v1.npz"))
    with DlaContext(k=20, num_samples=8) as ctx:
        for i in range(6):
            got = ctx.voigt(G["voigt_lambdas"], float(G["voigt_z"][i]), float(G["voigt_N"][i]), 3)
            assert np.max(np.abs(got - G["voigt_profiles"][i])) <= 1e-12
        got = ctx.voigt(G["voigt_lambdas"], float(G["voigt_z"][3]), 1e21, 31)
        assert np.max(np.abs(got - G["voigt_profile_31"])) <= 1e-12
    for name, Q, S, k, pre in (("tiny", 4, 64, 5, "tiny_"), ("config1", 2, 96, 20, "c1_")):
        params, model, samples, sp, prepared = prepared_config(name, Q, S)
        assert checksum(prepared) == str(G[pre + "checksum"])
        got = _run_gpu(prepared, samples, k, S, sp.log_priors_dla, sp.log_priors_no_dla)
        ref = {key[len(pre):]: G[key] for key in G.files if key.startswith(pre) and key != pre + "checksum"}
        assert_parity(got, ref)


@pytest.mark.parametrize("cfg,mask_gather", [("tiny", "reference"), ("config1", "reference"),
                                             ("config3", "reference"), ("config3", "intended")])
def test_device_side_preparation_matches_host_preparation(oracle, cfg, mask_gather):
    """dla_set_model + dla_eval_spectra (process_qsos.m:102-180 inside the library: selection on the
    host, interpolation / omega^2 scaling in a CUDA kernel) against the numpy preparation + dla_eval_batch
    and against the oracle on the numpy-prepared inputs."""
    Q, S = (4, 64) if cfg == "tiny" else (5, 128)
    params, model, samples, sp, prepared = prepared_config(cfg, Q, S, mask_gather)
    a = process_qsos(sp.all_wavelengths, sp.all_flux, sp.all_noise_variance, sp.all_pixel_mask, sp.z_qsos, model,
                     samples, sp.log_priors_dla, sp.log_priors_no_dla, params, mask_gather=mask_gather)
    b = process_qsos(sp.all_wavelengths, sp.all_flux, sp.all_noise_variance, sp.all_pixel_mask, sp.z_qsos, model,
                     samples, sp.log_priors_dla, sp.log_priors_no_dla, params, mask_gather=mask_gather,
                     device_prep=True)
    assert np.array_equal(a["min_z_dlas"], b["min_z_dlas"]) and np.array_equal(a["max_z_dlas"], b["max_z_dlas"])
    assert_parity(b, a)
    ref = oracle_batch(oracle, prepared, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    assert_parity(b, ref)
    rel = np.abs(a["sample_log_likelihoods_dla"] - b["sample_log_likelihoods_dla"]) / np.abs(a["sample_log_likelihoods_dla"])
    print(cfg, mask_gather, "device prep vs host prep, worst rel", rel.max())


@pytest.mark.parametrize("Q,n,k", [(5, 90, 5), (37, 1217, 20), (130, 333, 12)])
def test_training_objective_matches_the_objective_m_restatement(Q, n, k):
    """dla_train_objective (objective.m / spectrum_loss.m on the GPU: DMMA contraction over training quasars,
    per-quasar Cholesky + inverse, per-pixel gradient accumulation) against the numpy restatement."""
    from gp_dla_detection_b200.learn_qso_model import TrainingObjective
    from oracle import train_oracle as T
    x, y, l1, nv = T.make_training_set(Q, k=k, n_rest=n, seed=7 + Q)
    f_ref, g_ref = T.objective(x, y, l1, nv, apply_priors=True)
    obj = TrainingObjective(y, l1, nv, k)
    try:
        f, g = obj(x)
        f2, g2 = obj(x)
    finally:
        obj.close()
    assert abs(f - f_ref) <= 1e-10 * abs(f_ref)
    scale = np.max(np.abs(g_ref))
    assert np.max(np.abs(g - g_ref)) <= 1e-9 * scale, np.max(np.abs(g - g_ref)) / scale
    assert np.max(np.abs(g[-3:] - g_ref[-3:]) / np.abs(g_ref[-3:])) <= 1e-8
    assert f == f2 and np.array_equal(g, g2)          # fixed summation order: bitwise reproducible
    print("training objective: f rel err", abs(f - f_ref) / abs(f_ref), "g err/scale", np.max(np.abs(g - g_ref)) / scale)


def test_learn_qso_model_decreases_the_objective():
    from gp_dla_detection_b200.learn_qso_model import TrainingObjective, learn_qso_model
    from oracle import train_oracle as T
    x, y, l1, nv = T.make_training_set(64, k=4, n_rest=120, seed=3)
    obj = TrainingObjective(y, l1, nv, 4)
    f0 = obj(x)[0]
    obj.close()
    rest = 911.75 + 0.25 * np.arange(120)
    model, f1, res = learn_qso_model(rest, np.zeros(120), y, l1, nv, k=4, max_iter=30, x0=x)
    assert f1 < f0 and model.M.shape == (120, 4) and np.isfinite(model.log_beta)


def test_degenerate_shapes_do_not_crash(oracle):
    """Q = 0, S = 0 (only the no-DLA likelihood), S = 1, very short spectra (n < 16), NaN flux in one quasar."""
    params, model, samples, sp, prepared = prepared_config("tiny", 3, 8)
    k = prepared[0].M.shape[1]
    # Q = 0
    with DlaContext(k=k, num_samples=8) as ctx:
        ctx.set_samples(samples.offset_samples, samples.nhi_samples)
        out = ctx.eval_batch(PackedBatch([], k), np.zeros(0), np.zeros(0))
        assert out["log_likelihoods_dla"].shape == (0,)
    # S = 0: no hypotheses; the no-DLA likelihood is still evaluated
    with DlaContext(k=k, num_samples=0) as ctx:
        ctx.set_samples(np.zeros(0), np.zeros(0))
        out = ctx.eval_batch(PackedBatch(prepared, k), sp.log_priors_dla, sp.log_priors_no_dla)
    for q, p in enumerate(prepared):
        ref = oracle.log_mvnpdf_low_rank_c(p.flux, p.mu, p.M, p.omega2 + p.noise_variance)
        assert abs(out["log_likelihoods_no_dla"][q] - ref) <= 1e-10 * abs(ref)
    # S = 1 and a quasar cut down to 9 kept pixels
    import copy
    short = copy.deepcopy(prepared[0])
    short.n = 9
    for name in ("flux", "mu", "omega2", "noise_variance"):
        setattr(short, name, getattr(short, name)[:9].copy())
    short.M = np.asfortranarray(short.M[:9])
    short.profile_index = short.profile_index[:9].copy()
    off, nhi = samples.offset_samples[:1].copy(), samples.nhi_samples[:1].copy()
    with DlaContext(k=k, num_samples=1) as ctx:
        ctx.set_samples(off, nhi)
        got = ctx.eval_batch(PackedBatch([short, prepared[1]], k), sp.log_priors_dla[:2], sp.log_priors_no_dla[:2])

    class S1:
        offset_samples, nhi_samples = off, nhi
    ref = oracle_batch(oracle, [short, prepared[1]], S1, sp.log_priors_dla[:2], sp.log_priors_no_dla[:2])
    assert_parity(got, ref, check_map=False)
    assert list(got["map_ind"]) == [0, 0]
    # NaN flux: that quasar's likelihoods are NaN (as in MATLAB), the others are untouched
    bad = copy.deepcopy(prepared)
    bad[1].flux = bad[1].flux.copy()
    bad[1].flux[5] = np.nan
    a = _run_gpu(prepared, samples, k, 8, sp.log_priors_dla, sp.log_priors_no_dla)
    b = _run_gpu(bad, samples, k, 8, sp.log_priors_dla, sp.log_priors_no_dla)
    assert np.all(np.isnan(b["sample_log_likelihoods_dla"][1])) and np.isnan(b["log_likelihoods_no_dla"][1])
    assert b["map_ind"][1] == -1
    for q in (0, 2):
        assert np.array_equal(a["sample_log_likelihoods_dla"][q], b["sample_log_likelihoods_dla"][q])


def test_config4_dense_sampling_properties():
    """BASELINE config 4 shape (num_dla_samples = 100000) on two quasars: size-independent properties, and the
    first 10000 hypotheses of the dense set agree with a run on those 10000 alone (position in the sorted
    set only changes which series length a hypothesis shares with its warp)."""
    params, model, samples, sp, prepared = prepared_config("config4", 2, 100000)
    S = 100000
    got = _run_gpu(prepared, samples, 20, S, sp.log_priors_dla, sp.log_priors_no_dla)
    ll = got["sample_log_likelihoods_dla"]
    assert ll.shape == (2, S) and np.all(np.isfinite(ll))
    norm = np.exp(ll - (got["log_likelihoods_dla"][:, None] + np.log(S))).sum(axis=1)     # K10
    assert np.allclose(norm, 1.0, rtol=1e-10)
    assert np.array_equal(got["map_ind"], np.argmax(ll, axis=1))

    class Sub:
        offset_samples, nhi_samples = samples.offset_samples[:10000], samples.nhi_samples[:10000]
    sub = _run_gpu(prepared, Sub, 20, 10000, sp.log_priors_dla, sp.log_priors_no_dla)
    rel = np.abs(sub["sample_log_likelihoods_dla"] - ll[:, :10000]) / np.abs(ll[:, :10000])
    assert rel.max() <= 1e-10


def test_config1_full_sample_set_parity(oracle):
    """BASELINE config 1 at its full num_dla_samples = 10000 on six quasars (the C oracle with voigt.c verbatim
    takes ~0.5 s per quasar on 16 host threads): every north_star gate at the real size."""
    params, model, samples, sp, prepared = prepared_config("config1", 6, 10000)
    got = _run_gpu(prepared, samples, 20, 10000, sp.log_priors_dla, sp.log_priors_no_dla)
    ref = oracle_batch(oracle, prepared, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    worst = assert_parity(got, ref)
    print("config1 full-S worst per-sample rel err", worst,
          "log-posterior abs err", np.max(np.abs(got["log_posteriors_dla"] - ref["log_posteriors_dla"])))


def test_threaded_multi_context_sharding_matches_single_context():
    """process_qsos_multi_gpu: one host thread and one dla_ctx per shard (here all on cuda:0, or on two GPUs when
    the box has them); results are bitwise those of a single context, in batch order."""
    import torch
    from gp_dla_detection_b200.sharding import process_qsos_multi_gpu
    params, model, samples, sp, prepared = prepared_config("config3", 7, 128)
    args = (sp.all_wavelengths, sp.all_flux, sp.all_noise_variance, sp.all_pixel_mask, sp.z_qsos, model, samples,
            sp.log_priors_dla, sp.log_priors_no_dla, params)
    one = process_qsos(*args, device_prep=True)
    devs = (0, 1, 0) if torch.cuda.device_count() > 1 else (0, 0, 0)
    many = process_qsos_multi_gpu(*args, devices=devs)
    for key in ("sample_log_likelihoods_dla", "log_likelihoods_dla", "log_likelihoods_no_dla", "model_posteriors",
                "map_ind", "min_z_dlas", "max_z_dlas", "p_dlas", "map_z_dlas"):
        assert np.array_equal(one[key], many[key], equal_nan=True), key


@pytest.mark.parametrize("cfg,Q,S,k", [("config1", 100, 10000, 20), ("config3", 64, 2000, 20),
                                        ("config4", 2, 100000, 20), ("config5", 8, 1000, 40)])
def test_survey_parity_subsamples(oracle, cfg, Q, S, k):
    """parity on ALL of config 1 (100 quasars x 10000 hypotheses) and on seeded subsamples of the
    other configs, every north_star gate, against the C oracle (voigt.c verbatim when oracle/_ref is built; the
    oracle runs on all host threads, about a minute in total)."""
    params, model, samples, sp, prepared = prepared_config(cfg, Q, S)
    got = _run_gpu(prepared, samples, k, S, sp.log_priors_dla, sp.log_priors_no_dla)
    ref = oracle_batch(oracle, prepared, samples, sp.log_priors_dla, sp.log_priors_no_dla)
    worst = assert_parity(got, ref)
    dec = got["log_posteriors_dla"] > got["log_posteriors_no_dla"]
    print(f"{cfg}: {Q} quasars x {S}: worst per-sample rel err {worst:.2e}, log-posterior abs err "
          f"{np.max(np.abs(got['log_posteriors_dla'] - ref['log_posteriors_dla'])):.2e}, "
          f"{int(dec.sum())} DLA decisions (all equal to the oracle's)")


def test_eval_spectra_sub_batches_are_seamless():
    """dla_eval_spectra splits more than 2048 quasars into sub-batches (the host selection pass of one runs on a
    helper thread under the device work of the previous one): one call over 2100 quasars == the same quasars
    evaluated as [0, 2048) and [2048, 2100) in two calls, bit for bit, and == the host-prepared path."""
    from gp_dla_detection_b200 import synthetic
    from gp_dla_detection_b200.process_qsos import PackedSpectra
    Q, S = 2100, 64
    params, model, samples, sp = synthetic.make_config("tiny", num_quasars=Q, num_samples=S)
    pack = lambda a, b: PackedSpectra(sp.all_wavelengths[a:b], sp.all_flux[a:b], sp.all_noise_variance[a:b],
                                      sp.all_pixel_mask[a:b], sp.z_qsos[a:b])
    with DlaContext(k=5, num_lines=3, num_samples=S) as ctx:
        ctx.set_samples(samples.offset_samples, samples.nhi_samples)
        ctx.set_model(model, params)
        whole = ctx.eval_spectra(pack(0, Q), sp.log_priors_dla, sp.log_priors_no_dla)
        tm = ctx.timing()
        first = ctx.eval_spectra(pack(0, 2048), sp.log_priors_dla[:2048], sp.log_priors_no_dla[:2048])
        rest = ctx.eval_spectra(pack(2048, Q), sp.log_priors_dla[2048:], sp.log_priors_no_dla[2048:])
    assert tm["chunks"] >= 2
    for key in ("sample_log_likelihoods_dla", "log_likelihoods_dla", "log_likelihoods_no_dla", "log_posteriors_dla",
                "model_posteriors", "map_ind", "min_z_dlas", "max_z_dlas"):
        assert np.array_equal(whole[key], np.concatenate([first[key], rest[key]])), key
    sel = [0, 1, 2047, 2048, 2099]
    prepared = [prepare_quasar_(sp, q, model, params) for q in sel]
    host = _run_gpu(prepared, samples, 5, S, sp.log_priors_dla[sel], sp.log_priors_no_dla[sel])
    rel = np.abs(host["sample_log_likelihoods_dla"] - whole["sample_log_likelihoods_dla"][sel]) / \
        np.abs(host["sample_log_likelihoods_dla"])
    assert rel.max() <= 1e-10


def prepare_quasar_(sp, q, model, params):
    from gp_dla_detection_b200.process_qsos import prepare_quasar
    return prepare_quasar(sp.all_wavelengths[q], sp.all_flux[q], sp.all_noise_variance[q], sp.all_pixel_mask[q],
                          sp.z_qsos[q], model, params)


def test_eval_spectra_error_in_a_later_sub_batch_is_reported():
    """A quasar without any unmasked pixel in the modelled range is DLA_ERR_BAD_ARG (include/dla_b200.h) also when
    it sits in the second sub-batch, i.e. when the helper thread found it while the first sub-batch was running;
    the context stays usable afterwards.  Q = 0 is legal and returns empty arrays."""
    from gp_dla_detection_b200 import _lib, synthetic
    from gp_dla_detection_b200.process_qsos import PackedSpectra
    Q, S = 2100, 64
    params, model, samples, sp = synthetic.make_config("tiny", num_quasars=Q, num_samples=S)
    masks = [m.copy() for m in sp.all_pixel_mask]
    masks[2060][:] = True
    with DlaContext(k=5, num_lines=3, num_samples=S) as ctx:
        ctx.set_samples(samples.offset_samples, samples.nhi_samples)
        ctx.set_model(model, params)
        bad = PackedSpectra(sp.all_wavelengths, sp.all_flux, sp.all_noise_variance, masks, sp.z_qsos)
        with pytest.raises(_lib.DlaError) as ei:
            ctx.eval_spectra(bad, sp.log_priors_dla, sp.log_priors_no_dla)
        assert ei.value.status == -1 and "no unmasked pixel" in str(ei.value)
        good = PackedSpectra(sp.all_wavelengths[:8], sp.all_flux[:8], sp.all_noise_variance[:8],
                             sp.all_pixel_mask[:8], sp.z_qsos[:8])
        out = ctx.eval_spectra(good, sp.log_priors_dla[:8], sp.log_priors_no_dla[:8])
        assert np.all(np.isfinite(out["log_likelihoods_dla"]))
        empty = PackedSpectra([], [], [], [], np.zeros(0))
        out0 = ctx.eval_spectra(empty, np.zeros(0), np.zeros(0))
        assert out0["log_likelihoods_dla"].shape == (0,)
