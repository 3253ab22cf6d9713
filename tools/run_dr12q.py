"""BASELINE.json config 2 at FULL size, measured instead of projected: Q_total synthetic quasars (default 162 861,
the DR12Q count) x 10 000 DLA hypotheses, k = 20, 1217 pixels, sharded by quasar over the GPUs of one box.

    python tools/run_dr12q.py [--quasars 162861]                                   # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 tools/run_dr12q.py

Every rank generates its own shard of DISTINCT synthetic raw spectra (seeded by rank), keeps them in pinned host
memory, and makes ONE dla_eval_spectra call over the shard: raw spectra in, every process_qsos.m output out,
INCLUDING the full sample_log_likelihoods_dla matrix (13 GB over all ranks).  The timed region is the wall clock of
that call between barriers, max over ranks.  No collective on the data path; a side measurement, not bench.py's line.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quasars", type=int, default=162861)
    ap.add_argument("--samples", type=int, default=10000)
    ap.add_argument("--no-sample-ll", action="store_true", help="do not bring sample_log_likelihoods_dla back")
    ap.add_argument("--repeat", action="store_true", help="time a second identical call as well (buffers already sized)")
    ap.add_argument("--save", default=None, metavar="DIR",
                    help="write each rank's shard as DIR/processed_qsos_dr12q_rank<r>.mat (MATLAB v7.3 = HDF5, "
                         "process_qsos.m:245-249) through mat_io.save_processed_qsos and read one variable back")
    args = ap.parse_args()

    import torch
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("run_dr12q.py: no CUDA device; the hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(x, op):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=getattr(dist.ReduceOp, op))
        return float(t.item())

    from gp_dla_detection_b200 import synthetic
    from gp_dla_detection_b200.process_qsos import DlaContext, PackedSpectra
    from gp_dla_detection_b200.set_parameters import Parameters

    # equal shards (every quasar costs the same here: n = 1217 for all); the LPT partition of sharding.py reduces to this
    nq = args.quasars // world + (1 if rank < args.quasars % world else 0)
    S = args.samples
    params = Parameters(k=20, dlambda=0.25, num_dla_samples=S)
    seed = 20260002
    model = synthetic.make_model(20, 0.25, seed, params)
    samples = synthetic.make_samples(S, params)
    t0 = time.perf_counter()
    sp = synthetic.make_spectra(nq, model, seed + 1000 + 7919 * rank, False, params)
    t_gen = time.perf_counter() - t0

    keep = []

    def pin(a):
        t = torch.from_numpy(np.ascontiguousarray(a)).pin_memory()
        keep.append(t)
        return t.numpy()

    def pinned_alloc(shape, dtype):
        t = torch.empty(shape, dtype={np.float64: torch.float64, np.int32: torch.int32}[np.dtype(dtype).type]).pin_memory()
        keep.append(t)
        return t.numpy()

    nw = min(nq, 2048)   # warm-up slice: one sub-batch of the library
    warm = PackedSpectra(sp.all_wavelengths[:nw], sp.all_flux[:nw], sp.all_noise_variance[:nw],
                         sp.all_pixel_mask[:nw], sp.z_qsos[:nw])
    spectra = PackedSpectra(sp.all_wavelengths, sp.all_flux, sp.all_noise_variance, sp.all_pixel_mask, sp.z_qsos)
    sp.all_wavelengths = sp.all_flux = sp.all_noise_variance = sp.all_pixel_mask = None   # the packed copy is what is used
    for name in ("spec_ptr", "wavelengths", "flux", "noise_variance", "pixel_mask", "z_qso"):
        setattr(spectra, name, pin(getattr(spectra, name)))

    with DlaContext(k=20, num_lines=3, num_samples=S, device=local_rank, store_sample_ll=not args.no_sample_ll) as ctx:
        ctx.set_samples(samples.offset_samples, samples.nhi_samples)
        ctx.set_model(model, params)
        outputs = ctx._alloc_outputs(nq, alloc=pinned_alloc)
        # warm-up on one sub-batch of the library (2048 quasars: context, kernels, staging and device buffers at their
        # working size), then the one timed call over the whole shard
        ctx.eval_spectra(warm, sp.log_priors_dla[:nw], sp.log_priors_no_dla[:nw])
        barrier()
        t0 = time.perf_counter()
        out = ctx.eval_spectra(spectra, sp.log_priors_dla, sp.log_priors_no_dla, "reference", outputs)
        barrier()
        wall = time.perf_counter() - t0
        tm = ctx.timing()
        if args.repeat:
            barrier()
            t0 = time.perf_counter()
            out = ctx.eval_spectra(spectra, sp.log_priors_dla, sp.log_priors_no_dla, "reference", outputs)
            barrier()
            wall2 = time.perf_counter() - t0
        else:
            wall2 = float("nan")

    ok = bool(np.all(np.isfinite(out["log_likelihoods_dla"])) and np.all(np.isfinite(out["log_likelihoods_no_dla"])))
    if out["sample_log_likelihoods_dla"] is not None:
        ll = out["sample_log_likelihoods_dla"]
        ok = ok and bool(np.all(np.isfinite(ll[:: max(1, nq // 256)])))
        # K10 (calc_cddf.py:98, :244-246) on a strided subsample: normalised sample likelihoods sum to 1
        sub = slice(None, None, max(1, nq // 256))
        norm = np.exp(ll[sub] - (out["log_likelihoods_dla"][sub, None] + np.log(S))).sum(axis=1)
        ok = ok and bool(np.allclose(norm, 1.0, rtol=1e-10))
        ok = ok and bool(np.array_equal(out["map_ind"][sub], np.argmax(ll[sub], axis=1)))
    n_dla = int(np.sum(out["log_posteriors_dla"] > out["log_posteriors_no_dla"]))
    inj = int(np.sum(sp.injected))
    save_s = save_bytes = None
    if args.save:
        from gp_dla_detection_b200 import mat_io
        os.makedirs(args.save, exist_ok=True)
        path = os.path.join(args.save, f"processed_qsos_dr12q_rank{rank}.mat")
        res = dict(out)
        res["log_priors_dla"], res["log_priors_no_dla"] = sp.log_priors_dla, sp.log_priors_no_dla
        res["p_no_dlas"] = out["model_posteriors"][:, 0]                       # process_qsos.m:231
        res["p_dlas"] = 1 - res["p_no_dlas"]                                   # :232
        t0 = time.perf_counter()
        mat_io.save_processed_qsos(path, res, extra={"test_set_name": "dr12q_synthetic", "num_lines": 3,
                                                     "max_z_cut": params.max_z_cut, "min_z_cut": params.min_z_cut})
        save_s = time.perf_counter() - t0
        save_bytes = os.path.getsize(path)
        back = mat_io.load_processed_qsos(path)
        ok = ok and bool(np.array_equal(back["log_posteriors_dla"], out["log_posteriors_dla"]))
        if out["sample_log_likelihoods_dla"] is not None:
            ok = ok and bool(np.array_equal(back["sample_log_likelihoods_dla"][::97], out["sample_log_likelihoods_dla"][::97]))
        del back

    wall_max = reduce(wall, "MAX")
    total_q = int(round(reduce(float(nq), "SUM")))
    all_ok = reduce(1.0 if ok else 0.0, "MIN") == 1.0
    n_dla = int(round(reduce(float(n_dla), "SUM")))
    inj = int(round(reduce(float(inj), "SUM")))
    h2d = reduce(float(tm["h2d_bytes"]), "SUM")
    d2h = reduce(float(tm["d2h_bytes"]), "SUM")
    gen_max = reduce(t_gen, "MAX")
    wall2_max = reduce(wall2, "MAX") if args.repeat else None
    if args.save:
        save_max, save_total = reduce(save_s, "MAX"), reduce(float(save_bytes), "SUM")
    if rank == 0:
        print(json.dumps({
            "workload": "BASELINE config 2 at full size: %d distinct synthetic quasars x %d hypotheses, k=20, 1217 px" % (total_q, S),
            "n_gpus": world, "seconds": wall_max, "quasars_per_sec": total_q / wall_max,
            "timed": "wall clock of ONE dla_eval_spectra call per rank over its whole shard (raw spectra in pinned host "
                     "memory -> all outputs in pinned host memory), between barriers, max over ranks",
            "sample_log_likelihoods_returned": not args.no_sample_ll,
            "h2d_bytes": h2d, "d2h_bytes": d2h, "chunks_per_rank": int(tm["chunks"]),
            "device_ms_sum_over_chunks_rank0": {k: tm[k] for k in ("ms_total", "ms_voigt", "ms_contract", "ms_phi", "ms_reduce")},
            "second_call_seconds": wall2_max,
            "checks_passed": all_ok, "dla_decisions": n_dla, "injected_dlas": inj,
            "saved": ({"format": "MATLAB v7.3 (HDF5) via mat_io.save_processed_qsos / hdf5_min, one file per rank",
                       "seconds_max_over_ranks": save_max, "bytes_total": save_total, "read_back_equal": all_ok} if args.save else
                      "outputs not written to disk in this run (--save DIR writes processed_qsos v7.3 files)"),
            "generation_seconds_outside_timed_region": gen_max}))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
