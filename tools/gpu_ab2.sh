#!/bin/bash
# A/B on one box with environment variants: $1 = extra nvcc flags of build B, $2 = variants string for slices_diag
mkdir -p gpurun_out
run() { timeout 600 python tools/slices_diag.py --quasars 192 --oracle 0 --reps 4 --variants $2 2>&1 | grep '^{' | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l)
    if 'timing' in d and 'dmma' not in d['path']: print('$1', d['path'], {k: round(v,3) for k,v in d['timing'].items() if k in ('ms_total','ms_voigt','ms_contract','ms_chol')})"; }
for rep in 1 2; do
  DLA_NVCC_EXTRA="" python -m gp_dla_detection_b200.build --force > /dev/null 2>&1; run A "$2"
  DLA_NVCC_EXTRA="$1" python -m gp_dla_detection_b200.build --force > /dev/null 2>&1; run B "$2"
done
DLA_NVCC_EXTRA="" python -m gp_dla_detection_b200.build --force > /dev/null 2>&1
