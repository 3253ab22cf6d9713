#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_slices.py -m gpu -x -q > gpurun_out/l_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/l_pytest.log
timeout 600 python tools/slices_diag.py --quasars 192 --oracle 2 --reps 3 --variants DLA_B200_K2S_MODE=3 DLA_B200_K2S_MODE=2 DLA_B200_K2S_MODE=1 > gpurun_out/l_diag.log 2>&1; echo "diag rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/l_diag.log'):
    if l.startswith('{'):
        d=json.loads(l)
        if 'timing' in d: print(d['path'], {k: round(v,3) for k,v in d['timing'].items() if k.startswith('ms_')})
        else: print(d)
PY
