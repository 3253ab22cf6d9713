#!/bin/bash
# ncu --set full with source correlation of the Voigt slice kernel only (one launch, 16 quasars)
mkdir -p gpurun_out
F="python bench.py --steps 1 --warmup 1 --quasars 16 --no-cpu-baseline"
$F > gpurun_out/n_bench16.json 2> gpurun_out/n_bench16.err || { echo "plain failed"; tail -5 gpurun_out/n_bench16.err; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'voigt_slices' -c 1 \
    -f -o gpurun_out/r02_k1s $F > gpurun_out/n_ncu.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/n_ncu.log; ls -la gpurun_out/*.ncu-rep
