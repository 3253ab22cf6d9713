"""CPU suite: host-side mirror of process_qsos.m (steps 1-9), the C-ABI surface, and the
world_size-2 gloo path of the quasar sharding."""
import ctypes as C
import os
import re
import socket

import numpy as np
import pytest

from gp_dla_detection_b200 import _lib, synthetic
from gp_dla_detection_b200.process_qsos import PackedBatch, dla_existence_prior, prepare_quasar
from gp_dla_detection_b200.set_parameters import DEFAULT
from gp_dla_detection_b200.sharding import gather_outputs, partition_quasars

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_symbol_the_header_declares():
    hdr = open(os.path.join(ROOT, "include", "dla_b200.h")).read()
    declared = set(re.findall(r"\b(dla_[a-z_0-9]+)\s*\(", hdr))
    declared -= {"dla_ctx"}
    assert declared == set(_lib.EXPORTED_SYMBOLS)
    lib = _lib.load()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.dla_abi_version() == 2


def test_no_cpu_fallback_create_fails_loudly_without_a_device():
    lib = _lib.load()
    if lib.dla_device_count() > 0:
        pytest.skip("a CUDA device is present")
    from gp_dla_detection_b200.process_qsos import DlaContext
    with pytest.raises(_lib.DlaError) as ei:
        DlaContext()
    assert ei.value.status == -3 and "no CPU fallback" in str(ei.value)


def test_create_rejects_bad_parameters():
    lib = _lib.load()
    ctx = C.c_void_p()
    for k, nl in [(0, 3), (65, 3), (20, 0), (20, 32)]:  # k: 1..20 fused, 21..64 column-block path
        p = _lib.dla_params(k, nl, 10, 1, 0, 0)
        assert lib.dla_create(C.byref(ctx), 0, C.byref(p)) == -1
        assert lib.dla_last_error(None)


def test_prepare_quasar_follows_process_qsos_steps():
    params, model, samples, sp = synthetic.make_config("config3", num_quasars=3, num_samples=8)
    for q in range(3):
        lam, mask, z = sp.all_wavelengths[q], sp.all_pixel_mask[q].copy(), sp.z_qsos[q]
        mask[40:47] = True
        ref = prepare_quasar(lam, sp.all_flux[q], sp.all_noise_variance[q], mask, z, model, params, "reference")
        itd = prepare_quasar(lam, sp.all_flux[q], sp.all_noise_variance[q], mask, z, model, params, "intended")
        rest = lam / (1 + z)
        inr = (rest >= 911.75) & (rest <= 1215.75)
        keep = inr & ~mask
        assert ref.n_full == inr.sum() and ref.n == keep.sum() and ref.padded_wavelengths.shape[0] == ref.n_full + 6
        # K12: "reference" gathers the first n profile elements, "intended" the unmasked positions
        assert np.array_equal(ref.profile_index, np.arange(ref.n))
        assert np.array_equal(itd.profile_index, np.flatnonzero(~mask[inr]))
        assert np.array_equal(ref.flux, sp.all_flux[q][keep])
        # process_qsos.m:160-161, set_parameters.m:65-73
        kz = 3000 * 1000 / 299792458
        assert ref.max_z_dla == lam[keep].max() / 1215.6701 - 1 - kz
        assert ref.min_z_dla == max(lam[keep].min() / 1215.6701 - 1, 911.7633 * (1 + z) / 1215.6701 - 1 + kz)
        # process_qsos.m:169-177: three log-spaced pads each side, spacing 1e-4 dex
        pw = ref.padded_wavelengths
        assert np.array_equal(pw[3:-3], lam[inr])
        assert np.allclose(np.diff(np.log10(pw[:4])), 1e-4, rtol=0, atol=1e-12)
        assert np.allclose(np.diff(np.log10(pw[-4:])), 1e-4, rtol=0, atol=1e-12)
        # model interpolation + omega2 scaling (:139-147)
        lya_z = lam[keep] / 1215.6701 - 1
        sc = 1 - np.exp(-0.0023 * (1 + lya_z) ** 3.65) + 0.1
        om = np.exp(2 * np.interp(rest[keep], model.rest_wavelengths, model.log_omega)) * sc ** 2
        assert np.allclose(ref.omega2, om, rtol=1e-13)
        assert ref.M.shape == (ref.n, 20) and ref.M.flags.f_contiguous
    # unmasked case: both gather modes coincide
    a = prepare_quasar(sp.all_wavelengths[0], sp.all_flux[0], sp.all_noise_variance[0],
                       np.zeros_like(sp.all_pixel_mask[0]), sp.z_qsos[0], model, params, "intended")
    assert np.array_equal(a.profile_index, np.arange(a.n))


def test_prior_and_packing():
    zq = np.array([2.2, 2.5, 3.0, 3.5, 4.0])
    dla = np.array([0, 1, 0, 1, 0], dtype=bool)
    lp, ln = dla_existence_prior(3.0, zq, dla, DEFAULT)          # z < 3.0 + 0.1: first three
    assert lp == pytest.approx(np.log(1 / 3)) and ln == pytest.approx(np.log(2 / 3))
    lp, ln = dla_existence_prior(2.0, zq, np.zeros(5, bool), DEFAULT)
    assert lp == -np.inf or np.isnan(lp)                          # log(0) - log(0)/log(n): legal (process_qsos.m:129-132)
    params, model, samples, sp = synthetic.make_config("config3", num_quasars=4, num_samples=8)
    prep = [prepare_quasar(w, f, v, m, z, model, params) for w, f, v, m, z in
            zip(sp.all_wavelengths, sp.all_flux, sp.all_noise_variance, sp.all_pixel_mask, sp.z_qsos)]
    b = PackedBatch(prep, 20)
    assert b.pix_ptr[-1] == sum(p.n for p in prep) and b.pad_ptr[-1] == sum(p.n_full + 6 for p in prep)
    q = 2
    Mq = b.M[20 * b.pix_ptr[q]:20 * b.pix_ptr[q + 1]].reshape((prep[q].n, 20), order="F")
    assert np.array_equal(Mq, prep[q].M)


def test_partition_is_balanced_and_complete():
    rng = np.random.default_rng(0)
    costs = rng.integers(200, 1218, size=1000) * 10000.0
    shards = partition_quasars(costs, 8)
    allq = np.sort(np.concatenate(shards))
    assert np.array_equal(allq, np.arange(1000))
    loads = np.array([costs[s].sum() for s in shards])
    assert loads.max() / loads.mean() < 1.01


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    Q, S = 11, 5
    costs = np.arange(Q, dtype=float) + 1
    shards = partition_quasars(costs, world)
    mine = shards[rank]
    out = {"log_likelihoods_dla": mine * 10.0, "map_ind": mine.astype(np.int32),
           "sample_log_likelihoods_dla": np.outer(mine, np.ones(S)), "nothing": None}
    merged = gather_outputs(out, mine, Q, dist)
    dist.barrier()
    if rank == 0:
        q.put(merged)
    dist.destroy_process_group()


def test_world_size_2_gloo_host_gather():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    merged = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert np.array_equal(merged["log_likelihoods_dla"], np.arange(11) * 10.0)
    assert np.array_equal(merged["map_ind"], np.arange(11))
    assert np.array_equal(merged["sample_log_likelihoods_dla"][:, 0], np.arange(11.0))
    assert merged["nothing"] is None


def test_mex_shims_compile_and_link_against_the_c_abi(tmp_path):
    """No MATLAB/Octave here: compile the MEX gateways against the stand-in mex.h and link them with
    libdla_b200.so, proving they only use symbols the header declares."""
    import subprocess
    stub = tmp_path / "stub.c"
    stub.write_text('#include <stdarg.h>\n#include <stdlib.h>\n#include "mex.h"\n'
                    'void mexErrMsgIdAndTxt(const char *id, const char *fmt, ...) { (void)id; (void)fmt; abort(); }\n'
                    'int mexAtExit(void (*fn)(void)) { (void)fn; return 0; }\n')
    env = dict(os.environ); env.pop("CC", None)
    pkg = os.path.join(ROOT, "gp_dla_detection_b200")
    _lib.load()
    for src in ("voigt.c", "dla_eval_mex.c", "dla_process_mex.c"):
        out = tmp_path / (src + ".so")
        subprocess.run(["gcc", "-Wall", "-Werror", "-shared", "-fPIC", os.path.join(pkg, "mex", src), str(stub),
                        "-I", os.path.join(ROOT, "oracle", "shim"), "-I", os.path.join(ROOT, "include"),
                        "-L", pkg, "-l:libdla_b200.so", "-Wl,--no-undefined", "-o", str(out)], check=True, env=env)


def test_threaded_multi_gpu_helper_merges_in_batch_order_and_reraises(monkeypatch):
    """process_qsos_multi_gpu's host logic (LPT shards, one thread per device, merge back in batch order, worker
    exceptions re-raised on the caller's thread) with the per-shard evaluation replaced by a stand-in: no GPU."""
    import importlib
    from gp_dla_detection_b200 import sharding
    pq_mod = importlib.import_module("gp_dla_detection_b200.process_qsos")   # the package re-exports the function under this name
    Q, S = 23, 7
    rng = np.random.default_rng(1)
    lens = rng.integers(50, 90, size=Q)
    W = [np.full(n, float(i)) for i, n in enumerate(lens)]          # wavelengths[0] carries the quasar's identity
    seen = {}

    def fake_process_qsos(all_w, all_f, all_v, all_m, z, model, samples, lpd, lpn, params, device=0, **kw):
        ids = np.array([int(w[0]) for w in all_w])
        seen[device] = ids
        if kw.get("store_sample_ll") is False and device == 1:
            raise RuntimeError("boom on device 1")
        return {"log_likelihoods_dla": ids * 2.0 + np.asarray(lpd), "map_ind": ids.astype(np.int32),
                "sample_log_likelihoods_dla": np.outer(ids, np.ones(S)), "model_posteriors": np.tile(ids[:, None], (1, 2)) * 1.0,
                "nothing": None}

    monkeypatch.setattr(pq_mod, "process_qsos", fake_process_qsos)
    z = np.linspace(2.2, 4.0, Q)
    lpd = np.arange(Q) * 0.25
    out = sharding.process_qsos_multi_gpu(W, W, W, W, z, None, None, lpd, -1.0, devices=(0, 1, 2))
    assert np.array_equal(out["log_likelihoods_dla"], np.arange(Q) * 2.0 + lpd)
    assert np.array_equal(out["map_ind"], np.arange(Q))
    assert np.array_equal(out["sample_log_likelihoods_dla"][:, 3], np.arange(Q, dtype=float))
    assert out["nothing"] is None
    assert sorted(np.concatenate(list(seen.values())).tolist()) == list(range(Q)) and len(seen) == 3
    loads = [lens[ids].sum() for ids in seen.values()]
    assert max(loads) - min(loads) <= lens.max()                     # LPT balance by spectrum length
    with pytest.raises(RuntimeError, match="boom on device 1"):
        sharding.process_qsos_multi_gpu(W, W, W, W, z, None, None, lpd, -1.0, devices=(0, 1), store_sample_ll=False)
