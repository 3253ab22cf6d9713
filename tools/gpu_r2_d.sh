#!/bin/bash
mkdir -p gpurun_out
DLA_B200_TRACE=gpurun_out/trace.bin timeout 600 python tools/slices_diag.py --quasars 96 --samples 10000 --oracle 0 --reps 1 --only-default > gpurun_out/diag_trace.log 2>&1; echo "rc=$?"
cut -c1-400 gpurun_out/diag_trace.log | head -3
python tools/trace_report.py gpurun_out/trace.bin
