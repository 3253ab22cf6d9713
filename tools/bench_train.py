#!/usr/bin/env python
"""Side measurement (not the bench line): training-objective evaluations per second on one B200 against the
numpy restatement of objective.m on the host, on a synthetic training set of the reference's shape."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from gp_dla_detection_b200.learn_qso_model import TrainingObjective  # noqa: E402
from oracle import train_oracle as T  # noqa: E402

Q = int(sys.argv[1]) if len(sys.argv) > 1 else 5000
n, k = 1217, 20
x, y, l1, nv = T.make_training_set(Q, k=k, n_rest=n, seed=11)
obj = TrainingObjective(y, l1, nv, k)
f, g = obj(x)
t = time.perf_counter()
reps = 10
for _ in range(reps):
    f, g = obj(x)
dt = (time.perf_counter() - t) / reps
qs = min(Q, 40)
t = time.perf_counter()
fr, gr = T.objective(x, y[:qs], l1[:qs], nv[:qs], apply_priors=False)
dtc = (time.perf_counter() - t) / qs
fl, gl = TrainingObjective(y[:qs], l1[:qs], nv[:qs], k).local(x)
print(json.dumps({"what": "training objective (objective.m) f+g", "quasars": Q, "pixels": n, "k": k,
                  "gpu_ms_per_evaluation": 1e3 * dt, "gpu_quasars_per_s": Q / dt,
                  "cpu_numpy_quasars_per_s_1core": 1 / dtc, "sample": f"{qs} quasars on the host",
                  "f_rel_err_on_sample": abs(fl - fr) / abs(fr),
                  "g_err_over_scale_on_sample": float(np.max(np.abs(gl - gr)) / np.max(np.abs(gr)))}))
