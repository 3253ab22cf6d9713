#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_slices.py -m gpu -x -q > gpurun_out/k_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/k_pytest.log
timeout 600 python tools/parity_truth.py > gpurun_out/k_truth.log 2>&1; echo "truth rc=$?"
python - <<'PY'
import json
for r in json.load(open('gpurun_out/r02_parity_truth.json')):
    print(r['config'], {n: (r[n]['worst_abs'], r[n]['rms_abs']) for n in ('oracle_fp64','slices','dmma')})
PY
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/k_bench.json 2> gpurun_out/k_bench.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/k_bench.json')); print(d['value'], d['e2e']['value'], d['kernel_ms_per_step'])"
