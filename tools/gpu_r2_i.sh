#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_slices.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/i_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/i_pytest.log
timeout 600 python bench.py --workload config5 --quasars 16 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/i_cfg5.json 2> gpurun_out/i_cfg5.err; echo "cfg5 rc=$?"
python - <<'PY'
import json
try:
    d=json.loads([l for l in open('gpurun_out/i_cfg5.json') if l.startswith('{')][-1])
    print('value %.1f e2e %.1f'%(d['value'], d['e2e']['value']), d['kernel_ms_per_step'], d['roofline']['frac'])
except Exception as e:
    print('failed', e); print(open('gpurun_out/i_cfg5.err').read()[-2000:])
PY
