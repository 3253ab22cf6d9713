"""CPU: the host side of the training path -- rest-frame interpolation of learn_qso_model.m:29-75, the
pairwise-PCA initialisation, and the world_size-2 gloo all-reduce of the partial objectives (the numpy
restatement stands in for the per-rank CUDA evaluation, which needs a GPU)."""
import os
import socket

import numpy as np

from gp_dla_detection_b200 import synthetic
from gp_dla_detection_b200.learn_qso_model import add_priors, initial_x, prepare_training_set, reduce_over_ranks


def test_prepare_training_set_follows_learn_qso_model():
    params, model, samples, sp = synthetic.make_config("config3", num_quasars=6, num_samples=8)
    rest, mu, yc, l1, nv = prepare_training_set(sp.all_wavelengths, sp.all_flux, sp.all_noise_variance,
                                                sp.all_pixel_mask, sp.z_qsos, params, max_noise_variance=1.0)
    assert rest.shape == (1217,) and rest[0] == 911.75 and rest[-1] == 1215.75
    assert yc.shape == l1.shape == nv.shape == (6, 1217)
    ok = ~np.isnan(yc)
    assert ok.any() and not np.isnan(l1[ok]).any() and not np.isnan(nv[ok]).any()
    assert np.all(nv[ok] <= 1.0)                                   # :64-67
    with np.errstate(invalid="ignore"):
        assert np.allclose(np.nanmean(yc, axis=0)[ok.any(axis=0)], 0.0, atol=1e-12)   # centred (:70-71)
    q = 0
    z = sp.z_qsos[q]
    j = np.flatnonzero(ok[q])[10]
    assert abs(l1[q, j] - rest[j] * (1 + z) / 1215.6701) < 1e-9    # 1 + z_lya on the rest grid (:50-53)
    x0 = initial_x(yc, 5, params)
    assert x0.shape == (1217 * 6 + 3,) and np.all(np.isfinite(x0))
    assert np.allclose(x0[-3:], np.log([0.1, 0.0023, 3.65]))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    import torch.distributed as dist
    from oracle import train_oracle as T
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    x, y, l1, nv = T.make_training_set(9, k=3, n_rest=40, seed=5)
    mine = np.arange(9)[rank::world]                       # this rank's quasars
    f, g = T.objective(x, y[mine], l1[mine], nv[mine], apply_priors=False)
    f, g = reduce_over_ranks(f, g, dist)
    g = add_priors(x, g)
    dist.barrier()
    if rank == 0:
        q.put((f, g))
    dist.destroy_process_group()


def test_world_size_2_gloo_objective_allreduce():
    import torch.multiprocessing as mp
    from oracle import train_oracle as T
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    f, g = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    x, y, l1, nv = T.make_training_set(9, k=3, n_rest=40, seed=5)
    f1, g1 = T.objective(x, y, l1, nv, apply_priors=True)
    assert abs(f - f1) <= 1e-12 * abs(f1) and np.allclose(g, g1, rtol=1e-10, atol=1e-12)
