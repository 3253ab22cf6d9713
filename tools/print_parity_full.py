"""One line per case of gpurun_out/r02_parity_full.json."""
import json
import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for r in json.load(open(os.path.join(ROOT, "gpurun_out", "r02_parity_full.json"))):
    print(r["config"], r["mask_gather"], r.get("contraction"), "rel %.2e abs %.2e" % (r["worst_sample_ll_rel"], r["worst_sample_ll_abs"]),
          "pass", r["pass"], r.get("samples_beyond_relative_gate"))
