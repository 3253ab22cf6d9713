#!/bin/bash
# round 2: default bench line, the other BASELINE config shapes, the reference arm
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/h_bench_default.json 2> gpurun_out/h_bench_default.err; echo "default rc=$?"
for cfg in "config3 64" "config4 8" "config5 16"; do
  set -- $cfg
  timeout 600 python bench.py --workload $1 --quasars $2 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/h_cfg_$1.json 2> gpurun_out/h_cfg_$1.err
  echo "$1 rc=$?"
done
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/h_bench_reference.json 2> gpurun_out/h_bench_reference.err; echo "reference rc=$?"
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/h_*.json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][-1])
        k=d.get('kernel_ms_per_step',{})
        print(f, 'value %.1f e2e %.1f'%(d['value'], d['e2e']['value']), 'roof', d.get('roofline',{}).get('frac'), {a: round(b,3) for a,b in k.items()})
    except Exception as e:
        print(f, 'failed', e)
PY
