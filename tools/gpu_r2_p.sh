#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_slices.py -m gpu -x -q > gpurun_out/p_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/p_pytest.log
timeout 600 python tools/parity_truth.py > gpurun_out/p_truth.log 2>&1; echo "truth rc=$?"
python - <<'PY'
import json
for r in json.load(open('gpurun_out/r02_parity_truth.json')):
    print(r['config'], {n: (r[n]['worst_abs'], r[n]['rms_abs']) for n in ('oracle_fp64','slices','dmma')})
PY
bash tools/gpu_ab.sh "-DDLA_NSA=7"
