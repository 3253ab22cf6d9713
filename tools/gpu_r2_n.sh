#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_slices.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/n_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/n_pytest.log
timeout 600 python tools/slices_diag.py --quasars 192 --oracle 2 --reps 3 --variants DLA_B200_K1Q=1 > gpurun_out/n_diag.log 2>&1; echo "diag rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/n_diag.log'):
    if l.startswith('{'):
        d=json.loads(l)
        if 'timing' in d: print(d['path'], d.get('bitwise_equal_to_default'), {k: round(v,3) for k,v in d['timing'].items() if k in ('ms_total','ms_voigt','ms_contract','ms_chol')})
        else: print(d)
PY
timeout 600 python bench.py --workload config5 --quasars 16 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/n_cfg5.json 2> gpurun_out/n_cfg5.err; echo "cfg5 rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/n_cfg5.json')); print('config5', d['value'], d['kernel_ms_per_step'])"
