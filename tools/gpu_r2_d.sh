#!/bin/bash
mkdir -p gpurun_out
V="DLA_B200_K1Q=1 DLA_B200_K1Q=2 DLA_B200_K2S_STAGES=3 DLA_B200_OVERLAP=1"
timeout 600 python tools/slices_diag.py --quasars 192 --samples 10000 --oracle 2 --variants $V > gpurun_out/diag_var.log 2>&1; echo "rc=$?"
cut -c1-700 gpurun_out/diag_var.log
