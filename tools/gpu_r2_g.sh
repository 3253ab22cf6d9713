#!/bin/bash
# round 2, evidence run: plain bench (must exit 0), ncu launch list of the same command, ncu --set full of the slice kernels
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/g_bench_short.json 2> gpurun_out/g_bench_short.err || { echo "plain bench failed"; tail -5 gpurun_out/g_bench_short.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/g_launches.csv $B > gpurun_out/g_ncu_launch.log 2>&1
echo "launch list rc=$?"
F="python bench.py --steps 1 --warmup 1 --quasars 16 --no-cpu-baseline"
$F > gpurun_out/g_bench16.json 2> gpurun_out/g_bench16.err || { echo "plain bench16 failed"; tail -5 gpurun_out/g_bench16.err; exit 1; }
timeout 1200 ncu --set full --clock-control none --import-source on \
    -k regex:'voigt_slices|contract_slices|chol_rows|table_slices|table_colscale|pixel_scales|reduce|prep_model' -c 9 \
    -f -o gpurun_out/r02_slices $F > gpurun_out/g_ncu_full.log 2>&1
echo "ncu full rc=$?"; tail -3 gpurun_out/g_ncu_full.log; ls -la gpurun_out/*.ncu-rep
