#!/bin/bash
# usage: gpu_tune.sh VAR v1 v2 ...   -> one short bench per value of the environment variable
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
VAR=$1; shift
for v in "$@"; do
  env $VAR=$v timeout 300 python bench.py --steps 3 --warmup 3 --quasars 64 --no-cpu-baseline > gpurun_out/tune_$v.json 2> gpurun_out/tune_$v.err
  python - "$VAR=$v" gpurun_out/tune_$v.json <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[2])); k=d['kernel_ms_per_step']
    print(sys.argv[1], 'value %.0f e2e %.0f  voigt %.3f contract %.3f total %.3f'%(d['value'], d['e2e']['value'], k['ms_voigt'], k['ms_contract'], k['ms_total']))
except Exception as e:
    print(sys.argv[1], 'failed', e)
PY
done
