#!/bin/bash
# 1/2/4/8-GPU bench back to back on one box (as the driver's scaling step does)
mkdir -p gpurun_out
nvidia-smi -L | wc -l
for N in 1 2 4 8; do
  if [ $N = 1 ]; then
    timeout 600 python bench.py --gpus 1 --no-cpu-baseline > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+N)) \
      bench.py --gpus $N --no-cpu-baseline > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
  fi
  echo "N=$N rc=$?"
  python - gpurun_out/scale_n$N.json <<'PY'
import json,sys
try:
    line=[l for l in open(sys.argv[1]) if l.startswith('{')][-1]
    d=json.loads(line)
    print(d['n_gpus'], 'value %.0f e2e %.0f spectra %.0f ms/step %.2f'%(d['value'], d['e2e']['value'], d['e2e_from_spectra']['value'], d['ms_per_step']), d['clocks'])
except Exception as e:
    print('failed', e)
PY
done
