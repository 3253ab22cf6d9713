#!/bin/bash
# launch list + full ncu capture of the two hot kernels (no tests, no bench line)
mkdir -p gpurun_out
BENCH_SMALL="python bench.py --steps 1 --warmup 1 --quasars 16 --no-cpu-baseline"
timeout 300 $BENCH_SMALL > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $BENCH_SMALL > gpurun_out/ncu1.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'voigt_weights|contract_chol' -s 2 -c 2 -o gpurun_out/prof -f $BENCH_SMALL > gpurun_out/ncu2.log 2>&1
echo "ncu full rc=$?"
