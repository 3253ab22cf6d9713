#!/bin/bash
for b in 8 6 10 12; do
  DLA_NVCC_EXTRA="-DDLA_K1S_MINB=$b" python -m gp_dla_detection_b200.build --force > /dev/null 2>&1 || { echo "build failed $b"; continue; }
  grep -A3 "voigt_weights_lane_kernelE" gp_dla_detection_b200/build.log | grep -E "Used|spill" | tr '\n' ' '
  python bench.py --steps 3 --warmup 3 --quasars 64 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); k=d['kernel_ms_per_step']; print('MINB=$b', 'voigt %.3f total %.3f'%(k['ms_voigt'],k['ms_total']))"
done
