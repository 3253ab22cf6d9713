#!/bin/bash
# slice path: small diag (fails fast), then 64 and 192 quasars with the scheduling variants
mkdir -p gpurun_out
timeout 300 python tools/slices_diag.py --quasars 13 --samples 1000 --oracle 2 > gpurun_out/diag_small.log 2>&1; echo "diag_small rc=$?"
tail -6 gpurun_out/diag_small.log | cut -c1-600
V="DLA_B200_OVERLAP=0 DLA_B200_OVERLAP=0,DLA_B200_K2S_STAGES=3 DLA_B200_OVERLAP_CHUNKS=8 DLA_B200_OVERLAP_CHUNKS=32"
timeout 600 python tools/slices_diag.py --quasars 64 --samples 10000 --oracle 2 --variants $V > gpurun_out/diag_64.log 2>&1; echo "diag_64 rc=$?"
cut -c1-700 gpurun_out/diag_64.log
timeout 600 python tools/slices_diag.py --quasars 192 --samples 10000 --oracle 0 --variants $V > gpurun_out/diag_192.log 2>&1; echo "diag_192 rc=$?"
cut -c1-700 gpurun_out/diag_192.log
