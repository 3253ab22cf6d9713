#!/bin/bash
# ncu --set full of the slice-path kernels (K1s, K2s, K2c) on a 16-quasar config-2 batch
mkdir -p gpurun_out
CMD="python tools/slices_diag.py --quasars 16 --samples 10000 --oracle 0 --reps 1"
$CMD > gpurun_out/ncu_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'voigt_slices|contract_slices|chol_rows' -c 3 \
    -o gpurun_out/r02_slices $CMD > gpurun_out/ncu_run.log 2>&1
echo "ncu rc=$?"; tail -5 gpurun_out/ncu_run.log; ls -la gpurun_out/*.ncu-rep
