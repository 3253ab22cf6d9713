"""Quasar sharding across the GPUs of one box.

Every quasar is an independent unit, so the path shards by quasar with NO data-path collective:
a static longest-processing-time partition with cost n_q * S, each rank evaluates its shard on its
own GPU through its own dla_ctx, and the results are concatenated on the host ("final host gather").
torch.distributed is plumbing only (barrier, timing reduction, the host gather).
"""
from __future__ import annotations

import numpy as np


def partition_quasars(costs, world_size: int):
    """Static LPT partition. Returns a list (one entry per rank) of index arrays into the batch,
    each sorted ascending (stable output order inside a shard)."""
    costs = np.asarray(costs, dtype=np.float64)
    order = np.argsort(-costs, kind="stable")
    loads = np.zeros(world_size)
    shards = [[] for _ in range(world_size)]
    for q in order:
        r = int(np.argmin(loads))
        shards[r].append(int(q))
        loads[r] += costs[q]
    return [np.array(sorted(s), dtype=np.int64) for s in shards]


def gather_outputs(local_out: dict, local_index: np.ndarray, num_quasars: int, dist=None, dst: int = 0):
    """Concatenate per-rank output dicts (arrays with leading dimension = shard size) back into batch
    order on rank `dst`.  dist=None: single process (returns a reordered copy)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        pieces = [(local_index, local_out)]
    else:
        gathered = [None] * dist.get_world_size() if dist.get_rank() == dst else None
        dist.gather_object((local_index, local_out), gathered, dst=dst)
        if dist.get_rank() != dst:
            return None
        pieces = gathered
    merged = {}
    for key, val in pieces[0][1].items():
        if val is None:
            merged[key] = None
            continue
        proto = np.asarray(val)
        fill = np.nan if proto.dtype.kind == "f" else -1
        merged[key] = np.full((num_quasars,) + proto.shape[1:], fill, dtype=proto.dtype)
    for idx, out in pieces:
        for key, val in out.items():
            if val is not None and len(idx):
                merged[key][idx] = val
    return merged


def process_qsos_multi_gpu(all_wavelengths, all_flux, all_noise_variance, all_pixel_mask, z_qsos, model, samples,
                           log_priors_dla, log_priors_no_dla, params=None, devices=(0,), mask_gather="reference",
                           store_sample_ll=True, device_prep=True):
    """One process, several GPUs: the quasars are partitioned over ``devices`` (LPT by spectrum length), every
    shard is evaluated by its own host thread on its own dla_ctx (the C ABI allows one context per thread and
    device; ctypes releases the GIL during the calls), and the results are put back in batch order on the host.
    No collective anywhere.  ``devices`` may name a device twice (two contexts on one GPU) -- useful only for
    testing the plumbing."""
    import threading

    from .process_qsos import process_qsos
    from .set_parameters import DEFAULT

    params = params or DEFAULT
    Q = len(z_qsos)
    lpd = np.broadcast_to(np.asarray(log_priors_dla, dtype=np.float64), (Q,))
    lpn = np.broadcast_to(np.asarray(log_priors_no_dla, dtype=np.float64), (Q,))
    shards = partition_quasars([len(w) for w in all_wavelengths], len(devices))
    results, errors = [None] * len(devices), [None] * len(devices)

    def work(r):
        idx = shards[r]
        if len(idx) == 0:
            return
        try:
            pick = lambda xs: [xs[i] for i in idx]
            results[r] = process_qsos(pick(all_wavelengths), pick(all_flux), pick(all_noise_variance),
                                      pick(all_pixel_mask), np.asarray(z_qsos)[idx], model, samples, lpd[idx], lpn[idx],
                                      params, device=int(devices[r]), mask_gather=mask_gather,
                                      store_sample_ll=store_sample_ll, device_prep=device_prep)
        except BaseException as e:  # re-raised on the caller's thread
            errors[r] = e

    threads = [threading.Thread(target=work, args=(r,)) for r in range(len(devices))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for e in errors:
        if e is not None:
            raise e
    merged = None
    for r, out in enumerate(results):
        if out is None:
            continue
        piece = gather_outputs(out, shards[r], Q)
        if merged is None:
            merged = piece
        else:
            for key, val in piece.items():
                if val is not None:
                    merged[key][shards[r]] = val[shards[r]]
    return merged
