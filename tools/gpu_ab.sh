#!/bin/bash
# A/B of two builds on ONE box: $1 = extra nvcc flags of build B (build A = none); slices_diag on 192 quasars each, twice
mkdir -p gpurun_out
run() { timeout 600 python tools/slices_diag.py --quasars 192 --oracle 0 --reps 4 --only-default 2>&1 | grep '^{' | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print('$1', {k: round(v,3) for k,v in d['timing'].items() if k in ('ms_total','ms_voigt','ms_contract','ms_chol')})"; }
for rep in 1 2; do
  DLA_NVCC_EXTRA="" python -m gp_dla_detection_b200.build --force > /dev/null 2>&1; run A
  DLA_NVCC_EXTRA="$1" python -m gp_dla_detection_b200.build --force > /dev/null 2>&1; run "B($1)"
done
DLA_NVCC_EXTRA="" python -m gp_dla_detection_b200.build --force > /dev/null 2>&1
