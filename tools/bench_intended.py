#!/usr/bin/env python
"""Side measurement: config-3 shape with the masked ("intended") gather, lane kernel vs pixel-per-lane kernel
(DLA_B200_K1=pixel)."""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gp_dla_detection_b200.process_qsos import DlaContext, PackedBatch
from tests.helpers import prepared_config

params, model, samples, sp, prepared = prepared_config("config3", 32, 10000, "intended")
batch = PackedBatch(prepared, 20)
with DlaContext(k=20, num_samples=10000) as ctx:
    ctx.set_samples(samples.offset_samples, samples.nhi_samples)
    ctx.stage(batch, sp.log_priors_dla, sp.log_priors_no_dla)
    for _ in range(2):
        ctx.run_staged()
    ms = {"ms_voigt": 0.0, "ms_contract": 0.0, "ms_total": 0.0}
    for _ in range(3):
        ctx.run_staged()
        t = ctx.timing()
        for k in ms:
            ms[k] += t[k] / 3
    out = ctx.fetch()
print(json.dumps({"K1": os.environ.get("DLA_B200_K1", "lane"), "quasars": 32, **ms,
                  "checksum": float(np.nansum(out["log_likelihoods_dla"]))}))
