"""Where does the worst per-sample disagreement between the CUDA path and the C oracle come from?
(diagnostic, run on the GPU box: python tools/diag_parity.py)"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tests.helpers import prepared_config, oracle_batch
from tests.test_gpu_parity import _run_gpu
from oracle import oracle as O

Q = int(sys.argv[1]) if len(sys.argv) > 1 else 100
params, model, samples, sp, prepared = prepared_config("config1", Q, 10000)
got = _run_gpu(prepared, samples, 20, 10000, sp.log_priors_dla, sp.log_priors_no_dla)
a = got["sample_log_likelihoods_dla"]
for use_ref in (True, False):
    ref = oracle_batch(O, prepared, samples, sp.log_priors_dla, sp.log_priors_no_dla, use_ref_voigt=use_ref)
    b = ref["sample_log_likelihoods_dla"]
    err = np.abs(a - b); rel = err / np.abs(b)
    print("oracle voigt =", "voigt.c verbatim" if use_ref else "restated", "| worst rel", rel.max(), "worst abs", err.max(),
          "| median abs", np.median(err), "| 99.9% abs", np.quantile(err, 0.999))
    idx = np.argsort(rel.ravel())[::-1][:8]
    for i in idx:
        q, s = divmod(int(i), a.shape[1])
        z = prepared[q].min_z_dla + (prepared[q].max_z_dla - prepared[q].min_z_dla) * samples.offset_samples[s]
        print(f"  q={q} s={s} ll={b[q, s]:.6f} abs={err[q, s]:.3e} rel={rel[q, s]:.3e} z={z:.5f} logN={np.log10(samples.nhi_samples[s]):.3f}")
    big = err > 1e-8
    print("  samples with abs err > 1e-8:", int(big.sum()), "of", big.size)
