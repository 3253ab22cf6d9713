"""Full-size parity subsamples SURVEY.md 8(d) promises, too slow for `pytest -m gpu`:

    config 3 (ragged 200-1217 px)        64 quasars x 10 000 hypotheses, both gather modes
    config 5 (k = 40, 2433 px)           16 quasars x 10 000
    config 4 (num_dla_samples = 100 000)  8 quasars x 100 000
    config 1                             100 quasars x 10 000 (all of it)

CUDA path through the C ABI against the C oracle (voigt.c verbatim where oracle/_ref is built), every north_star
gate; worst errors to gpurun_out/r02_parity_full.json (copied to profiles/ by hand).  Run on the GPU box:

    python tools/parity_full.py [--quick]
"""
from __future__ import annotations

import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from gp_dla_detection_b200.process_qsos import DlaContext, PackedBatch  # noqa: E402
from tests.helpers import ATOL_LOG_POST, RTOL_SAMPLE_LL, assert_parity, oracle_batch, prepared_config  # noqa: E402


def main():
    quick = "--quick" in sys.argv
    from oracle import oracle as O
    O.load_c()
    cases = [("config1", 100, 10000, 20, "reference"), ("config3", 64, 10000, 20, "reference"),
             ("config3", 64, 10000, 20, "intended"), ("config5", 16, 10000, 40, "reference"),
             ("config4", 8, 100000, 20, "reference")]
    if quick:
        cases = [(c, max(2, Q // 16), S // 10, k, g) for c, Q, S, k, g in cases]
    rows = []
    for cfg, Q, S, k, gather in cases:
        t0 = time.time()
        params, model, samples, sp, prepared = prepared_config(cfg, Q, S, gather)
        with DlaContext(k=k, num_samples=S) as ctx:
            ctx.set_samples(samples.offset_samples, samples.nhi_samples)
            got = ctx.eval_batch(PackedBatch(prepared, k), sp.log_priors_dla, sp.log_priors_no_dla)
            got["_contraction"] = ctx.timing()["contraction"]
        t1 = time.time()
        ref = oracle_batch(O, prepared, samples, sp.log_priors_dla, sp.log_priors_no_dla)
        t2 = time.time()
        a, b = got["sample_log_likelihoods_dla"], ref["sample_log_likelihoods_dla"]
        rel = np.abs(a - b) / np.abs(b)
        ok, why = True, ""
        try:
            assert_parity(got, ref)
        except AssertionError as e:   # keep going: the record must show every case
            ok, why = False, str(e)
        # samples beyond the relative gate: where do the two fp64 results stand against the x87 extended-precision
        # value of the reference's formula (oracle/lmlr_extended.c, from the oracle's own fp64 profile)?
        beyond = np.argwhere(rel > RTOL_SAMPLE_LL)
        voigt = O.voigt_ref if O.load_ref() is not None else O.voigt_c
        bg = {"count": int(beyond.shape[0]), "checked": 0, "abs_ll_max": None, "cuda_minus_exact_worst": None,
              "oracle_minus_exact_worst": None, "cuda_closer_than_oracle": 0}
        for q, s_i in beyond[:100]:
            pq = prepared[q]
            zz = pq.min_z_dla + (pq.max_z_dla - pq.min_z_dla) * samples.offset_samples[s_i]
            prof = voigt(pq.padded_wavelengths, zz, samples.nhi_samples[s_i], 3)[pq.profile_index]
            exact = O.log_mvnpdf_low_rank_extended(pq.flux, pq.mu * prof, np.asarray(pq.M) * prof[:, None],
                                                   pq.omega2 * prof ** 2 + pq.noise_variance)
            ec, eo = abs(a[q, s_i] - exact), abs(b[q, s_i] - exact)
            bg["checked"] += 1
            bg["abs_ll_max"] = max(bg["abs_ll_max"] or 0.0, abs(float(b[q, s_i])))
            bg["cuda_minus_exact_worst"] = max(bg["cuda_minus_exact_worst"] or 0.0, float(ec))
            bg["oracle_minus_exact_worst"] = max(bg["oracle_minus_exact_worst"] or 0.0, float(eo))
            bg["cuda_closer_than_oracle"] += int(ec <= eo)
        dec = got["log_posteriors_dla"] > got["log_posteriors_no_dla"]
        dec_ref = ref["log_posteriors_dla"] > ref["log_posteriors_no_dla"]
        row = {"config": cfg, "mask_gather": gather, "quasars": Q, "num_dla_samples": S, "k": k,
               "pixels_min_max": [int(min(p.n for p in prepared)), int(max(p.n for p in prepared))],
               "worst_sample_ll_rel": float(np.nanmax(rel)), "median_sample_ll_rel": float(np.nanmedian(rel)),
               "worst_sample_ll_abs": float(np.nanmax(np.abs(a - b))),
               "worst_log_posterior_abs": float(np.max(np.abs(got["log_posteriors_dla"] - ref["log_posteriors_dla"]))),
               "worst_ll_no_dla_abs": float(np.max(np.abs(got["log_likelihoods_no_dla"] - ref["log_likelihoods_no_dla"]))),
               "map_index_equal": int(np.sum(got["map_ind"] == ref["map_ind"])),
               "decisions_equal": int(np.sum(dec == dec_ref)), "dla_decisions": int(dec.sum()),
               "gates": {"sample_ll_rel": RTOL_SAMPLE_LL, "log_posterior_abs": ATOL_LOG_POST},
               "pass": ok, "failure": why, "samples_beyond_relative_gate": bg,
               "contraction": "int8 slices" if got.get("_contraction") == 1 else "fp64 DMMA", "oracle": "voigt.c verbatim + C restatement" if O.load_ref() is not None
               else "C restatement", "gpu_s": round(t1 - t0, 2), "oracle_s": round(t2 - t1, 2)}
        rows.append(row)
        print(json.dumps(row), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "r02_parity_full.json"), "w") as f:
        json.dump(rows, f, indent=1)
    sys.exit(0 if all(r["pass"] for r in rows) else 1)


if __name__ == "__main__":
    main()
