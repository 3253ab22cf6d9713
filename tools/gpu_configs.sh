#!/bin/bash
# side measurements: throughput at the other BASELINE.json config shapes (not bench lines)
mkdir -p gpurun_out
for cfg in "config3 64" "config4 8" "config5 16"; do
  set -- $cfg
  timeout 600 python bench.py --workload $1 --quasars $2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/cfg_$1.json 2> gpurun_out/cfg_$1.err
  python - $1 gpurun_out/cfg_$1.json <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[2])); k=d['kernel_ms_per_step']
    print(sys.argv[1], 'value %.1f q/s e2e %.1f  K2 frac %.3f step frac %.3f voigt %.3f contract %.3f total %.3f'%(d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['step_frac_end_to_end'], k['ms_voigt'], k['ms_contract'], k['ms_total']))
except Exception as e:
    print(sys.argv[1], 'failed', e); print(open(sys.argv[2][:-4]+'err').read()[-1500:])
PY
done
