#!/bin/bash
# rebuild on the GPU box with different K2 stage geometries and time the bench step
for cfg in "8 4" "16 2" "4 8"; do
  set -- $cfg
  DLA_NVCC_EXTRA="-DDLA_K2P_KC=$1 -DDLA_K2P_STAGES=$2" python -m gp_dla_detection_b200.build --force > /dev/null 2>&1 || { echo "build failed $cfg"; continue; }
  python bench.py --steps 3 --warmup 3 --quasars 64 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['kernel_ms_per_step']; print('KC=$1 STAGES=$2', 'contract %.3f total %.3f'%(k['ms_contract'],k['ms_total']))"
done
