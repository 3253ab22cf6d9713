"""How far is the REFERENCE'S OWN fp64 arithmetic from the exact value of its formula, and how far are the CUDA paths?

For samples of a config (the worst |CUDA - oracle| ones and a random set) the Gaussian log-pdf of
log_mvnpdf_low_rank.m is evaluated once more in x87 extended precision (oracle/lmlr_extended.c, 64-bit mantissa)
from the oracle's own fp64 absorption profile, and three fp64 results are compared with it: the oracle (= the
reference's statements in fp64), the int8 slice path and the FP64 DMMA path.  Worst and rms absolute differences to
gpurun_out/r02_parity_truth.json.       python tools/parity_truth.py   (GPU box; test infrastructure)
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from gp_dla_detection_b200.process_qsos import DlaContext, PackedBatch  # noqa: E402
from tests.helpers import oracle_batch, prepared_config  # noqa: E402


def gpu(env, prepared, samples, sp, k, S):
    for key in ("DLA_B200_CONTRACT",):
        os.environ.pop(key, None)
    os.environ.update(env)
    with DlaContext(k=k, num_samples=S) as ctx:
        ctx.set_samples(samples.offset_samples, samples.nhi_samples)
        out = ctx.eval_batch(PackedBatch(prepared, k), sp.log_priors_dla, sp.log_priors_no_dla)
        assert ctx.timing()["contraction"] == (0 if env else 1)
    os.environ.pop("DLA_B200_CONTRACT", None)
    return out["sample_log_likelihoods_dla"]


def main():
    from oracle import oracle as O
    O.load_c()
    voigt = O.voigt_ref if O.load_ref() is not None else O.voigt_c
    rows = []
    for cfg, Q, S, k, nworst, nrand in (("config1", 6, 10000, 20, 60, 300), ("config3", 8, 10000, 20, 60, 300),
                                         ("config5", 3, 10000, 40, 40, 150)):
        params, model, samples, sp, prepared = prepared_config(cfg, Q, S, "reference")
        a_sl = gpu({}, prepared, samples, sp, k, S)
        a_dm = gpu({"DLA_B200_CONTRACT": "dmma"}, prepared, samples, sp, k, S)
        ref = oracle_batch(O, prepared, samples, sp.log_priors_dla, sp.log_priors_no_dla)["sample_log_likelihoods_dla"]
        rng = np.random.default_rng(11)
        err = {"oracle_fp64": [], "slices": [], "dmma": []}
        small = {"oracle_fp64": [], "slices": [], "dmma": []}
        for q, p in enumerate(prepared):
            worst = np.argsort(-np.abs(a_sl[q] - ref[q]))[:nworst]
            near0 = np.argsort(np.abs(ref[q]))[:20]                   # the samples a relative gate is hardest on
            pick = np.unique(np.concatenate([worst, near0, rng.choice(S, nrand, replace=False)]))
            z = p.min_z_dla + (p.max_z_dla - p.min_z_dla) * samples.offset_samples[pick]      # process_qsos.m:163-165
            for s_i, zz in zip(pick, z):
                a = voigt(p.padded_wavelengths, zz, samples.nhi_samples[s_i], 3)[p.profile_index]   # :186-189
                truth = O.log_mvnpdf_low_rank_extended(p.flux, p.mu * a, np.asarray(p.M) * a[:, None],
                                                       p.omega2 * a ** 2 + p.noise_variance)         # :191-197
                for name, v in (("oracle_fp64", ref[q, s_i]), ("slices", a_sl[q, s_i]), ("dmma", a_dm[q, s_i])):
                    err[name].append(v - truth)
                    if abs(truth) < 1.0:
                        small[name].append((v - truth) / truth)
        row = {"config": cfg, "quasars": Q, "num_dla_samples": S, "k": k, "samples_checked": len(err["slices"]),
               "truth": "log_mvnpdf_low_rank.m in x87 extended precision from the oracle's fp64 profile",
               "typical_abs_ll": float(np.median(np.abs(ref)))}
        for name in err:
            e = np.asarray(err[name])
            row[name] = {"worst_abs": float(np.max(np.abs(e))), "rms_abs": float(np.sqrt(np.mean(e * e))),
                         "worst_rel_where_abs_ll_below_1": float(np.max(np.abs(small[name]))) if small[name] else None}
        row["samples_with_abs_ll_below_1"] = len(small["slices"])
        rows.append(row)
        print(json.dumps(row), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(rows, open(os.path.join(ROOT, "gpurun_out", "r02_parity_truth.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
