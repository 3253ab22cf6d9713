#!/usr/bin/env python
"""bench.py -- quasars/sec of the DLA model-evaluation hot path on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One step = one pass of the hot path (K0 tables, K1 Voigt+weights, K2 contraction + Cholesky, K3
log-mean-exp/argmax) over one batch of synthetic quasars of BASELINE.json config-2 shape (1217
pixels, k = 20, S = 10000 hypotheses).  For k <= 20 the contraction is the exact-slice int8 tcgen05
kernel (DLA_B200_CONTRACT=dmma selects the FP64 DMMA kernels instead; k > 20 always uses them).  Quasars shard across ranks with
no collective on the data path (weak scaling: --quasars per GPU per step).  Rank 0 prints ONE
JSON line.  `value` is device-resident throughput (inputs already in HBM), `e2e` goes through
dla_eval_batch with pinned HOST buffers (H2D + D2H inside the timed region).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "quasars/sec"
N_PIX, K_RANK = 1217, 20
# K2's DRAM traffic per quasar from the committed ncu --set full capture (3.2578 GB read + 0.0061 GB written
# by one 16-quasar launch, profiles/r01_ncu_full_summary.txt); not measured inside bench.py
NCU_K2_DRAM_GB_PER_QUASAR = (3.257835 + 0.006098) / 16
# the same for the int8 slice contraction (profiles/r02_ncu_slices_summary.txt: 2.919 GB read + 0.325 GB written by one
# 16-quasar launch) and the FP64 instructions of the slice-emitting Voigt kernel per (hypothesis, pixel)
NCU_K2S_DRAM_GB_PER_QUASAR = (2.469627 + 0.321808) / 16   # profiles/r02_ncu_slices_summary.txt
NCU_K1S_FP64_PER_PIXEL = 56.84   # DFMA + DMUL + DADD warp instructions x 32 / (hypothesis, pixel) pairs (ncu source page by opcode)
SLICE_PAIRS = 27                 # sample slices i < 6 x table slices j < 7 with i + j <= 6 (csrc/slices_path.cuh, NSA = 6)
PLANE_BYTES_PER_PAIR = 12.0      # int8 planes of W and U per (hypothesis, pixel): 2 x 6 slices
INT8_NOMINAL_TOPS = 4500.0       # dense 8-bit tensor peak NVIDIA quotes for B200 (= fp8)
FP64_PIPE_TFLOPS = 37.1   # register-resident DMMA / DFMA loops, tools/probe_fp64.cu, profiles/r01_probe_fp64.jsonl
# BASELINE.json configs as (rank, grid spacing, ragged masking, hypotheses); SURVEY.md 8(d)
WORKLOADS = {
    "config2": {"k": 20, "dlambda": 0.25, "ragged": False, "samples": 10000, "pixels": 1217,
                "name": "BASELINE config 2 shape (DR12Q-scale synthetic batch)"},
    "config3": {"k": 20, "dlambda": 0.25, "ragged": True, "samples": 10000, "pixels": 1217,
                "name": "BASELINE config 3 shape (ragged realistic masking, 200-1217 kept pixels)"},
    "config4": {"k": 20, "dlambda": 0.25, "ragged": False, "samples": 100000, "pixels": 1217,
                "name": "BASELINE config 4 shape (dense prior sampling, 100000 hypotheses)"},
    "config5": {"k": 40, "dlambda": 0.125, "ragged": False, "samples": 10000, "pixels": 2433,
                "name": "BASELINE config 5 shape (stressed contraction, k=40 on the 0.125 A grid)"},
    # side measurement of the training objective (objective.m); --quasars = training quasars
    "train": {"k": 20, "dlambda": 0.25, "ragged": False, "samples": 0, "pixels": 1217,
              "name": "training objective f+g (objective.m / spectrum_loss.m)"},
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--quasars", type=int, default=192, help="quasars per GPU per step")
    ap.add_argument("--samples", type=int, default=None, help="num_dla_samples (config 2: 10000)")
    ap.add_argument("--workload", default="config2", choices=sorted(WORKLOADS),
                    help="BASELINE.json config shape; config2 is the bench line, the others are side measurements")
    ap.add_argument("--cpu-baseline-seconds", type=float, default=2.0,
                    help="wall seconds of the CPU-baseline sample (x host threads = CPU work, ~10-30 core-seconds)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    global N_PIX, K_RANK
    w = WORKLOADS[args.workload]
    K_RANK, N_PIX = w["k"], w["pixels"]
    if args.samples is None:
        args.samples = w["samples"]
    return args


# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi clocks/throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.gpu), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def measure_dgemm_tflops(torch, n=8192, reps=5):
    """FP64 tensor roofline denominator: cuBLAS DGEMM n^3, best of reps (MEASURED_PEAKS.json has none)."""
    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    b = torch.randn(n, n, dtype=torch.float64, device="cuda")
    torch.matmul(a, b)
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); torch.matmul(a, b); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    del a, b
    torch.cuda.empty_cache()
    return 2.0 * n ** 3 / (best * 1e-3) / 1e12


# ------------------------------------------------------------------------------------------------
def cpu_reference_rate(prepared, samples, seconds: float, nthreads: int = 0):
    """Time the CPU implementation of the path on a bounded sample of the same workload:
    voigt.c compiled verbatim (oracle/_ref) for the profile + the plain-C restatement of
    process_qsos.m:184-209 / log_mvnpdf_low_rank.m as written, OpenMP over samples.
    The sample grows quasar by quasar until about `seconds` of wall time is spent."""
    from oracle import oracle as O
    lib = O.load_c()
    have_ref = O.load_ref() is not None
    # all host cores this process may use (torchrun exports OMP_NUM_THREADS=1: do not rely on it)
    cores = nthreads if nthreads > 0 else max(lib.oracle_max_threads(), len(os.sched_getaffinity(0)))
    S = samples.offset_samples.shape[0]
    lp = (np.log(0.1), np.log(0.9))
    n0 = min(S, 64 * cores)
    t = time.perf_counter()
    O.process_quasar_c(prepared[0], samples.offset_samples, samples.nhi_samples, *lp, use_ref_voigt=have_ref,
                       nthreads=cores, first_sample=0, num_eval=n0)
    dt0 = time.perf_counter() - t
    n1 = int(min(S, max(n0, n0 * seconds / max(dt0, 1e-3))))   # hypotheses per timed call
    done, dt, nq = 0, 0.0, 0
    while dt < seconds and nq < len(prepared):
        t = time.perf_counter()
        O.process_quasar_c(prepared[nq], samples.offset_samples, samples.nhi_samples, *lp, use_ref_voigt=have_ref,
                           nthreads=cores, first_sample=0, num_eval=n1)
        dt += time.perf_counter() - t
        done += n1
        nq += 1
        if n1 < S:
            break
    return {"value": (done / S) / dt, "unit": METRIC, "cores": cores,
            "kind": "port",
            "sample": f"{n1} of {S} hypotheses of {nq} quasar(s) ({N_PIX} px, k={K_RANK}) in {dt:.2f}s "
                      f"(= {dt * cores:.0f} core-seconds); "
                      f"profile = {'reference voigt.c compiled verbatim' if have_ref else 'C restatement of voigt.c'}"
                      f" + libcerf stand-in; log_mvnpdf_low_rank.m as written in C; OpenMP over samples"}, n1, dt


def slice_gcols(k: int) -> int:
    """Row length of G on the slice path (slice_layout, csrc/slices_path.cuh): the W.Phi and U.M segments in units of 64
    columns, a rest of <= 32 columns in a unit of 32."""
    if k <= 20:
        return 256
    def seg(c):
        r = c % 64
        return 64 * (c // 64 + (1 if r > 32 else 0)) + (32 if 0 < r <= 32 else 0)
    return seg(k * (k + 1) // 2) + seg(k)


def make_workload(args, rank):
    from gp_dla_detection_b200 import synthetic
    from gp_dla_detection_b200.process_qsos import PackedBatch, prepare_quasar
    from gp_dla_detection_b200.set_parameters import Parameters
    params = Parameters(k=K_RANK, num_dla_samples=args.samples)
    w = WORKLOADS[args.workload]
    model = synthetic.make_model(K_RANK, w["dlambda"], 20260002, params)
    samples = synthetic.make_samples(args.samples, params)
    sp = synthetic.make_spectra(args.quasars, model, 20260002 + 1000 + rank, w["ragged"], params)
    prepared = [prepare_quasar(w, f, v, m, z, model, params)
                for w, f, v, m, z in zip(sp.all_wavelengths, sp.all_flux, sp.all_noise_variance, sp.all_pixel_mask,
                                         sp.z_qsos)]
    make_workload.model = model
    return params, samples, sp, prepared, PackedBatch(prepared, K_RANK)


def config_dict(args, world):
    return {"workload": f"{WORKLOADS[args.workload]['name']}: {N_PIX} px, k={K_RANK}, "
                        f"S={args.samples} hypotheses/quasar; {args.quasars} quasars per GPU per step, sharded by "
                        f"quasar over {world} GPU(s), no collective; 162861 quasars = "
                        f"{162861 / (args.quasars * world):.0f} such steps",
            "quasars_per_gpu_per_step": args.quasars, "num_dla_samples": args.samples, "pixels": N_PIX, "k": K_RANK,
            "l2": "per-step working set (W,U weight tiles, %.1f GB/GPU) far exceeds the 126 MB L2"
                  % (args.quasars * 2 * 8 * (args.samples + 64) * (N_PIX + 15) / 1e9)}


# ------------------------------------------------------------------------------------------------
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = args.gpus
    params, samples, sp, prepared, batch = make_workload(argparse.Namespace(**{**vars(args), "quasars": 1}), 0)
    per_step = max(2.0, min(30.0, 100.0 / max(1, args.steps + args.warmup)))
    info, n1, _ = cpu_reference_rate(prepared, samples, per_step)
    from oracle import oracle as O
    have_ref = O.load_ref() is not None
    S = args.samples
    times = []
    for it in range(args.warmup + args.steps):
        t = time.perf_counter()
        O.process_quasar_c(prepared[0], samples.offset_samples, samples.nhi_samples, np.log(0.1), np.log(0.9),
                           use_ref_voigt=have_ref, nthreads=info["cores"], first_sample=0, num_eval=n1)
        if it >= args.warmup:
            times.append(time.perf_counter() - t)
    total = sum(times)
    value = (n1 / S) * args.steps / total
    info["value"] = value
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": METRIC, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": config_dict(args, world), "cpu_baseline": info,
            "e2e": {"value": value, "unit": METRIC, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "reference under Octave not runnable (no Octave/mkoctfile/libcerf); this arm times the "
                    "reference's CPU algorithm: voigt.c verbatim + C restatement of the .m files, all host threads"}
    print(json.dumps(line), flush=True)


def run_ours(args):
    import torch
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"   # keep NCCL's version banner off stdout: rank 0 prints ONE JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    from gp_dla_detection_b200.process_qsos import DlaContext
    params, samples, sp, prepared, batch = make_workload(args, rank)
    S, Q = args.samples, batch.Q

    # pinned host copies of the step inputs for the e2e leg
    def pin(a):
        t = torch.from_numpy(a).pin_memory()
        return t.numpy(), t
    keep = []
    for name in ("pix_ptr", "pad_ptr", "padded_wavelengths", "profile_index", "flux", "mu", "omega2", "noise_variance",
                 "M", "min_z_dla", "max_z_dla"):
        arr, t = pin(getattr(batch, name))
        setattr(batch, name, arr)
        keep.append(t)

    dgemm_tflops = measure_dgemm_tflops(torch)

    ctx = DlaContext(k=K_RANK, num_lines=3, num_samples=S, device=local_rank, store_sample_ll=True)
    ctx.set_samples(samples.offset_samples, samples.nhi_samples)
    ctx.stage(batch, sp.log_priors_dla, sp.log_priors_no_dla)

    # ---- device-resident leg -----------------------------------------------------------------
    for _ in range(args.warmup):
        ctx.run_staged()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    t0 = time.perf_counter()
    ev_ms = {"ms_total": 0.0, "ms_phi": 0.0, "ms_voigt": 0.0, "ms_contract": 0.0, "ms_chol": 0.0, "ms_reduce": 0.0}
    launches = 0
    for _ in range(args.steps):
        ctx.run_staged()
        tm = ctx.timing()
        for k_ in ev_ms:
            ev_ms[k_] += tm[k_]
        launches += tm["launches"]
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop()
    dev_s = max_over_ranks(ev_ms["ms_total"] * 1e-3)   # CUDA events on the library's stream, max over ranks
    wall_s = max_over_ranks(wall)
    value = Q * world * args.steps / dev_s

    # ---- end-to-end leg: dla_eval_batch, pinned host buffers in, host buffers out ---------------
    def pinned_alloc(shape, dtype):
        t = torch.empty(shape, dtype={np.float64: torch.float64, np.int32: torch.int32}[np.dtype(dtype).type]).pin_memory()
        keep.append(t)
        return t.numpy()
    outputs = ctx._alloc_outputs(Q, alloc=pinned_alloc)
    ctx.eval_batch(batch, sp.log_priors_dla, sp.log_priors_no_dla, outputs)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out = ctx.eval_batch(batch, sp.log_priors_dla, sp.log_priors_no_dla, outputs)
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    tm = ctx.timing()
    e2e = {"value": Q * world * args.steps / e2e_s, "unit": METRIC, "h2d_bytes_per_step": int(tm["h2d_bytes"]),
           "d2h_bytes_per_step": int(tm["d2h_bytes"]), "timed": "wall clock around K dla_eval_batch calls (pinned host buffers in and out; copies of one quasar chunk overlap the kernels of another), max over ranks",
           "host_prep": "per-quasar numpy preparation (process_qsos.m:102-180) is outside the timed region"}
    assert np.all(np.isfinite(out["log_likelihoods_dla"]))

    # ---- end to end from RAW spectra: per-quasar preparation (process_qsos.m:102-180) inside the library ----
    from gp_dla_detection_b200.process_qsos import PackedSpectra
    spectra = PackedSpectra(sp.all_wavelengths, sp.all_flux, sp.all_noise_variance, sp.all_pixel_mask, sp.z_qsos)
    for name in ("spec_ptr", "wavelengths", "flux", "noise_variance", "pixel_mask", "z_qso"):
        arr, t = pin(getattr(spectra, name))
        setattr(spectra, name, arr)
        keep.append(t)
    ctx.set_model(make_workload.model, params)
    out2 = ctx.eval_spectra(spectra, sp.log_priors_dla, sp.log_priors_no_dla, "reference", outputs)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        out2 = ctx.eval_spectra(spectra, sp.log_priors_dla, sp.log_priors_no_dla, "reference", outputs)
    barrier()
    raw_s = max_over_ranks(time.perf_counter() - t0)
    tm2 = ctx.timing()
    e2e_raw = {"value": Q * world * args.steps / raw_s, "unit": METRIC, "h2d_bytes_per_step": int(tm2["h2d_bytes"]),
               "d2h_bytes_per_step": int(tm2["d2h_bytes"]),
               "what": "dla_eval_spectra: raw ragged spectra (wavelengths, flux, noise variance, pixel mask, z_qso) in pinned "
                       "host memory -> results; selection / padded grid on the host inside the library, model "
                       "interpolation and omega^2 scaling on the device (learned model resident)"}
    assert np.allclose(out2["log_likelihoods_dla"], out["log_likelihoods_dla"], rtol=1e-9)

    # ---- one quasar per call: the reference's own loop shape (process_qsos.m:88 `for quasar_ind = 1:num_quasars`)
    # with the per-spectrum MEX gateway: dla_set_samples + dla_set_model (both return at once when the arrays are
    # the ones already resident) + dla_eval_spectra(Q = 1), host arrays in, host results out, every call.
    nq1 = min(Q, 48)
    singles = [PackedSpectra(sp.all_wavelengths[q:q + 1], sp.all_flux[q:q + 1], sp.all_noise_variance[q:q + 1],
                             sp.all_pixel_mask[q:q + 1], sp.z_qsos[q:q + 1]) for q in range(nq1)]
    out1 = ctx._alloc_outputs(1, alloc=pinned_alloc)

    def one_by_one():
        acc = 0.0
        for q in range(nq1):
            ctx.set_samples(samples.offset_samples, samples.nhi_samples)
            ctx.set_model(make_workload.model, params)
            o = ctx.eval_spectra(singles[q], sp.log_priors_dla[q:q + 1], sp.log_priors_no_dla[q:q + 1], "reference", out1)
            acc += float(o["log_likelihoods_dla"][0])
        return acc
    one_by_one()
    barrier()
    t0 = time.perf_counter()
    acc1 = one_by_one()
    barrier()
    q1_s = max_over_ranks(time.perf_counter() - t0)
    tm1 = ctx.timing()
    e2e_q1 = {"value": nq1 * world / q1_s, "unit": METRIC, "quasars_timed": nq1, "ms_per_call": 1e3 * q1_s / nq1,
              "device_ms_per_call": tm1["ms_total"], "h2d_bytes_per_call": int(tm1["h2d_bytes"]),
              "d2h_bytes_per_call": int(tm1["d2h_bytes"]),
              "what": "Q = 1 per call, the shape of process_qsos.m's own loop: dla_set_samples + dla_set_model (unchanged "
                      "arrays: two memcmp's) + dla_eval_spectra of ONE quasar, wall clock per call"}
    assert np.isfinite(acc1) and abs(acc1 - float(out2["log_likelihoods_dla"][:nq1].sum())) <= 1e-6 * abs(acc1)

    # ---- rooflines ------------------------------------------------------------------------------
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    n_sum = float(batch.n.sum())
    flop_per_step = S * n_sum * K_RANK * (K_RANK + 3)            # n k (k+3) per (quasar, sample)
    t_contract = ev_ms["ms_contract"] * 1e-3 / args.steps
    t_voigt = ev_ms["ms_voigt"] * 1e-3 / args.steps
    ach = flop_per_step / t_contract / 1e12
    k2_name = ("contract_chol_persistent_kernel (K2: persistent DMMA contraction + Cholesky epilogue)" if K_RANK <= 20 else
               "contract_block_kernel + chol_generic_kernel (K2 generic rank: column-block DMMA contraction, smem Cholesky)")
    slices = int(tm.get("contraction", 0)) == 1
    step_tflops = flop_per_step / (dev_s / args.steps) / 1e12
    if not slices:
        roofline = {"kernel": k2_name, "bound": "tensor",
                    "achieved": ach, "peak": dgemm_tflops, "unit": "TFLOP/s", "frac": ach / dgemm_tflops,
                    "peak_pipe": FP64_PIPE_TFLOPS,
                    "frac_of_pipe": ach / FP64_PIPE_TFLOPS,
                    "traffic": (NCU_K2_DRAM_GB_PER_QUASAR * Q / max(1, tm["chunks"])) if args.workload == "config2" and S == 10000 else None,
                    "traffic_unit": "GB per launch (dram__bytes_read.sum + dram__bytes_write.sum of profiles/r01_ncu_full_summary.txt, "
                                    "scaled from its 16-quasar launch; algorithmic W+U bytes are 0.198 GB per quasar)",
                    "peak_source": "FP64 tensor peak = cuBLAS DGEMM 8192^3 measured in this run (MEASURED_PEAKS.json has no FP64 "
                                   "figure); peak_pipe = register-resident DMMA loop of tools/probe_fp64.cu "
                                   "(profiles/r01_probe_fp64.jsonl), a constant, not re-measured here",
                    "algorithmic_flop_per_launch": flop_per_step / max(1, tm["chunks"]),
                    "launches_per_step": int(tm["chunks"]), "ms_per_step": 1e3 * t_contract,
                    "step_frac_end_to_end": step_tflops / dgemm_tflops}
    else:
        # int8 slice contraction: SLICE_PAIRS slice pairs x (224 + 32) columns x n_pad32 pixels x S_pad128 samples, 2 ops per MAC;
        # the tensor peak for 8-bit operands is twice the measured bf16 figure (same pipe, K = 32 instead of 16 per MMA)
        spad = (S + 1 + 127) // 128 * 128
        npad32 = float(((batch.n + 31) // 32 * 32).sum())
        gcols = slice_gcols(K_RANK)
        int8_ops = 2.0 * SLICE_PAIRS * gcols * spad * npad32
        # denominator: the burst figure (the sustained one is what cuBLAS bf16 reaches at ITS power draw, 1.24 GHz here;
        # the int8 contraction draws less per operation and runs at 1.7 GHz, so it can exceed twice the sustained figure)
        bf16_peak = float(peaks.get("bf16_tflops", 1650.0))
        int8_peak = 2.0 * bf16_peak
        int8_sustained = 2.0 * float(peaks.get("bf16_tflops_sustained", 1400.0))
        ach8 = int8_ops / t_contract / 1e12
        roofline = {"kernel": "contract_slices_tiles_kernel (K2s: exact-slice int8 tcgen05 contraction, 6 x 7 slices, "
                              "int32 TMEM accumulators, fp64 recombination)",
                    "bound": "tensor", "achieved": ach8, "peak": int8_peak, "unit": "TFLOP/s", "frac": ach8 / int8_peak,
                    "unit_note": "tensor-core operations per second on 8-bit integer operands (one multiply-add = 2), i.e. int8 TOP/s; "
                                 "the fp64-equivalent flop rate is in fp64_equivalent",
                    "peak_nominal": INT8_NOMINAL_TOPS, "frac_of_nominal": ach8 / INT8_NOMINAL_TOPS,
                    "peak_sustained": int8_sustained, "frac_of_sustained": ach8 / int8_sustained, "g_columns": gcols,
                    "shape_bound": "the 7 x 7 slice scheme issues MMAs of 32..256 columns; measured cost max(N/2, 64 + N/6) clocks each "
                                   "(profiles/r02_probe_umma.jsonl) caps a 64-column unit at 0.855 and a 32-column unit at 0.63 of the pipe's "
                                   "peak rate: 0.78 for the 3+2 unit mix; ncu: tensor pipe active 64 % of cycles",
                    "peak_source": ("2 x MEASURED_PEAKS.json bf16_tflops (burst; 8-bit operands run at twice the bf16 rate on the "
                                    "same pipe: tools/probe_umma.cu); peak_sustained = 2 x bf16_tflops_sustained"
                                    if peaks else "2 x fallback 1650 / 1400 TFLOP/s bf16 burst / sustained"),
                    "algorithmic_int8_op_per_launch": int8_ops / max(1, tm["chunks"]),
                    "fp64_equivalent": {"achieved_tflops": ach, "dgemm_peak_tflops": dgemm_tflops, "ratio": ach / dgemm_tflops,
                                        "note": "n k (k+3) fp64 flops per (quasar, hypothesis) (SURVEY 8d) / K2s time, against the "
                                                "cuBLAS DGEMM 8192^3 measured in this run: the contraction no longer runs on the FP64 pipe"},
                    "traffic": (NCU_K2S_DRAM_GB_PER_QUASAR * Q / max(1, tm["chunks"])) if args.workload == "config2" and S == 10000 and K_RANK == 20 else None,
                    "traffic_unit": "GB per launch (dram__bytes_read.sum + dram__bytes_write.sum of profiles/r02_ncu_slices_summary.txt, "
                                    "scaled from its 16-quasar launch; algorithmic bytes: int8 planes of W and U 0.151 GB + G 0.021 GB per quasar)",
                    "launches_per_step": int(tm["chunks"]), "ms_per_step": 1e3 * t_contract,
                    "step_fp64_equivalent_vs_dgemm": step_tflops / dgemm_tflops}
    if slices:
        # K1s is bound by the FP64 vector pipe: FP64 instructions per (hypothesis, pixel) from the ncu capture of the same
        # kernel (sm__inst_executed_pipe_fp64 / (S_pad n_pad)), as DFMA-equivalent TFLOP/s against the FP64 peak
        spad = (S + 1 + 127) // 128 * 128
        npad32 = float(((batch.n + 31) // 32 * 32).sum())
        fl = 2.0 * NCU_K1S_FP64_PER_PIXEL * spad * npad32
        roofline_voigt = {"kernel": "voigt_slices_lane_kernel (K1s, sample-per-lane, emits int8 slice planes)",
                          "bound": "fp64 vector pipe", "achieved": fl / t_voigt / 1e12, "peak": dgemm_tflops,
                          "peak_pipe": FP64_PIPE_TFLOPS, "unit": "TFLOP/s (2 x FP64 instructions)",
                          "frac": fl / t_voigt / 1e12 / dgemm_tflops, "frac_of_pipe": fl / t_voigt / 1e12 / FP64_PIPE_TFLOPS,
                          "fp64_instructions_per_hypothesis_pixel": NCU_K1S_FP64_PER_PIXEL,
                          "plane_write_gbs": PLANE_BYTES_PER_PAIR * spad * npad32 / t_voigt / 1e9, "hbm_peak_gbs": hbm_peak,
                          "faddeeva_evals_per_s": 3.0 * S * float((batch.n_full + 6).sum()) / t_voigt,
                          "ms_per_step": 1e3 * t_voigt,
                          "peak_source": "cuBLAS DGEMM measured in this run; peak_pipe = DFMA/DMMA probe constant "
                                         "(profiles/r01_probe_fp64.jsonl); instruction count from profiles/r02_ncu_slices_summary.txt"}
    else:
        wu_bytes = 16.0 * S * n_sum                                    # W and U written once (fp64)
        roofline_voigt = {"kernel": "voigt_weights_lane_kernel (K1, sample-per-lane)", "bound": "fp64 vector pipe",
                          "achieved": 2.0 * 58 * S * n_sum / t_voigt / 1e12, "peak": dgemm_tflops, "peak_pipe": FP64_PIPE_TFLOPS,
                          "unit": "TFLOP/s (2 x FP64 instructions, 58 per hypothesis and pixel: profiles/r01_ncu_full_summary.txt)",
                          "frac": 2.0 * 58 * S * n_sum / t_voigt / 1e12 / dgemm_tflops,
                          "wu_write_gbs": wu_bytes / t_voigt / 1e9, "hbm_peak_gbs": hbm_peak,
                          "faddeeva_evals_per_s": 3.0 * S * float((batch.n_full + 6).sum()) / t_voigt,
                          "ms_per_step": 1e3 * t_voigt}
    if slices:
        dominant = "K1s voigt_slices_lane_kernel" if t_voigt > t_contract else "K2s contract_slices_tiles_kernel"
    else:
        dominant = "K1 voigt_weights_lane_kernel" if t_voigt > t_contract else "K2 " + k2_name.split(" ")[0]

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            try:
                cpu, _, _ = cpu_reference_rate(prepared, samples, args.cpu_baseline_seconds)
            except Exception as e:  # the checker is optional for the product's own number
                cpu = {"value": None, "unit": METRIC, "cores": 0, "kind": "port", "sample": f"failed: {e}"}
        line = {"metric": METRIC, "value": value, "unit": METRIC, "n_gpus": world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": 1e3 * dev_s / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": config_dict(args, world), "e2e": e2e, "e2e_from_spectra": e2e_raw, "e2e_q1": e2e_q1,
                "gpu_launches": int(launches),
                "clocks": clocks, "roofline": roofline, "roofline_voigt": roofline_voigt,
                "dominant_kernel": dominant, "cpu_baseline": cpu,
                "kernel_ms_per_step": {k_: v / args.steps for k_, v in ev_ms.items()},
                "wall_ms_per_step": 1e3 * wall_s / args.steps,
                "dr12q_projection_s": 162861 / value}
        print(json.dumps(line), flush=True)
    ctx.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def run_train(args):
    """Side measurement, not the bench line: evaluations of the training objective on one B200 against the
    numpy restatement of objective.m (the checker) on a bounded sample of the same training set."""
    from gp_dla_detection_b200 import synthetic
    from gp_dla_detection_b200.learn_qso_model import TrainingObjective
    Q = args.quasars
    x, y, l1, nv = synthetic.make_training_set(Q, k=K_RANK, n_rest=N_PIX, seed=11)
    obj = TrainingObjective(y, l1, nv, K_RANK)
    for _ in range(max(1, args.warmup)):
        f, g = obj(x)
    t = time.perf_counter()
    for _ in range(args.steps):
        f, g = obj(x)
    dt = (time.perf_counter() - t) / args.steps
    cpu = None
    if not args.no_cpu_baseline:
        from oracle import train_oracle as T
        qs = min(Q, 40)
        t = time.perf_counter()
        fr, gr = T.objective(x, y[:qs], l1[:qs], nv[:qs], apply_priors=False)
        dtc = time.perf_counter() - t
        sub = TrainingObjective(y[:qs], l1[:qs], nv[:qs], K_RANK)
        fl, gl = sub.local(x)
        sub.close()
        cpu = {"value": qs / dtc, "unit": "training quasars/sec", "cores": 1, "kind": "port",
               "sample": f"{qs} quasars, numpy restatement of objective.m",
               "f_rel_err_on_sample": abs(fl - fr) / abs(fr),
               "g_err_over_scale_on_sample": float(np.max(np.abs(gl - gr)) / np.max(np.abs(gr)))}
    obj.close()
    print(json.dumps({"metric": "training quasars/sec", "value": Q / dt, "unit": "training quasars/sec", "n_gpus": 1,
                      "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt, "higher_is_better": True,
                      "dtype": "f64", "data": "synthetic",
                      "config": {"workload": WORKLOADS["train"]["name"], "quasars": Q, "pixels": N_PIX, "k": K_RANK},
                      "timed": "wall clock around dla_train_objective (x H2D, f/g D2H inside)", "cpu_baseline": cpu}),
          flush=True)


def main():
    args = parse_args()
    if args.workload == "train":
        return run_train(args)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
