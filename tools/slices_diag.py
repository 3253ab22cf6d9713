"""Slices path against the DMMA path and the C oracle on the same inputs, plus per-kernel device times.

    python tools/slices_diag.py [--quasars 64] [--samples 10000] [--oracle 2] [--modes]

Prints one JSON object per line.  --modes also times the contraction's diagnosis builds (DLA_B200_K2S_MODE=1: no MMA,
2: no loads after the first ring fill); their results are not valid likelihoods.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from gp_dla_detection_b200.process_qsos import DlaContext, PackedBatch  # noqa: E402
from tests.helpers import oracle_batch, prepared_config  # noqa: E402


def run(env, prepared, samples, sp, k, S, reps):
    for key in ("DLA_B200_CONTRACT", "DLA_B200_K2S_MODE", "DLA_B200_OVERLAP", "DLA_B200_K2S_STAGES", "DLA_B200_OVERLAP_CHUNKS"):
        os.environ.pop(key, None)
    os.environ.update(env)
    with DlaContext(k=k, num_samples=S) as ctx:
        ctx.set_samples(samples.offset_samples, samples.nhi_samples)
        batch = PackedBatch(prepared, k)
        ctx.stage(batch, sp.log_priors_dla, sp.log_priors_no_dla)
        ctx.run_staged()
        best = None
        for _ in range(reps):
            ctx.run_staged()
            tm = ctx.timing()
            if best is None or tm["ms_total"] < best["ms_total"]:
                best = tm
        out = ctx.fetch()
    return out, best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quasars", type=int, default=64)
    ap.add_argument("--samples", type=int, default=10000)
    ap.add_argument("--oracle", type=int, default=2)
    ap.add_argument("--config", default="config1")
    ap.add_argument("--modes", action="store_true")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--only-default", action="store_true", help="one context with the environment as it is (tracing)")
    ap.add_argument("--variants", nargs="*", default=[], help="extra runs, each a comma list of ENV=value")
    a = ap.parse_args()
    params, model, samples, sp, prepared = prepared_config(a.config, a.quasars, a.samples, "reference")
    k = 20
    t0 = time.time()
    out_s, tm_s = run({}, prepared, samples, sp, k, a.samples, a.reps)
    if a.only_default:
        print(json.dumps({"path": "slices (overlapped chunks)", "timing": tm_s}), flush=True)
        return
    out_d, tm_d = run({"DLA_B200_CONTRACT": "dmma"}, prepared, samples, sp, k, a.samples, a.reps)
    print(json.dumps({"path": "slices (overlapped chunks)", "timing": tm_s}), flush=True)
    print(json.dumps({"path": "dmma", "timing": tm_d}), flush=True)
    for env in a.variants:
        e = dict(kv.split("=") for kv in env.split(",") if kv)
        o, tm = run(e, prepared, samples, sp, k, a.samples, a.reps)
        same = bool(np.array_equal(o["sample_log_likelihoods_dla"], out_s["sample_log_likelihoods_dla"], equal_nan=True))
        print(json.dumps({"path": "slices " + env, "bitwise_equal_to_default": same, "timing": tm}), flush=True)
    x, y = out_s["sample_log_likelihoods_dla"], out_d["sample_log_likelihoods_dla"]
    rel = np.abs(x - y) / np.abs(y)
    print(json.dumps({"compare": "slices vs dmma", "quasars": a.quasars, "nan_slices": int(np.isnan(x).sum()),
                      "nan_dmma": int(np.isnan(y).sum()), "worst_rel": float(np.nanmax(rel)),
                      "worst_abs": float(np.nanmax(np.abs(x - y))), "median_rel": float(np.nanmedian(rel)),
                      "ll_no_abs": float(np.max(np.abs(out_s["log_likelihoods_no_dla"] - out_d["log_likelihoods_no_dla"]))),
                      "map_equal": int(np.sum(out_s["map_ind"] == out_d["map_ind"])),
                      "gpu_s": round(time.time() - t0, 1)}), flush=True)
    if a.oracle > 0:
        from oracle import oracle as O
        O.load_c()
        nq = min(a.oracle, a.quasars)
        ref = oracle_batch(O, prepared[:nq], samples, sp.log_priors_dla[:nq], sp.log_priors_no_dla[:nq])
        r = ref["sample_log_likelihoods_dla"]
        for name, o in (("slices", out_s), ("dmma", out_d)):
            g = o["sample_log_likelihoods_dla"][:nq]
            print(json.dumps({"compare": f"{name} vs oracle", "quasars": nq,
                              "worst_rel": float(np.nanmax(np.abs(g - r) / np.abs(r))),
                              "worst_abs": float(np.nanmax(np.abs(g - r))),
                              "median_abs": float(np.nanmedian(np.abs(g - r))),
                              "ll_no_abs": float(np.max(np.abs(o["log_likelihoods_no_dla"][:nq] - ref["log_likelihoods_no_dla"])))}),
                  flush=True)
    if a.modes:
        for mode in ("1", "2"):
            _, tm = run({"DLA_B200_K2S_MODE": mode}, prepared, samples, sp, k, a.samples, a.reps)
            print(json.dumps({"path": f"slices, K2S_MODE={mode}", "timing": tm}), flush=True)


if __name__ == "__main__":
    main()
