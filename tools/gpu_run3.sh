#!/bin/bash
# gpu tests, bench, launch list + full ncu capture of the two hot kernels
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu.log; echo "pytest rc=${PIPESTATUS[0]}"
tail -8 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 3 --warmup 3 --quasars 64 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print({k:d[k] for k in ('value','ms_per_step')}, d['e2e']['value'], d['roofline']['frac'], d['kernel_ms_per_step'])
PY
tail -5 gpurun_out/bench.err
if [ "$1" != "noncu" ]; then
BENCH_SMALL="python bench.py --steps 1 --warmup 1 --quasars 16 --no-cpu-baseline"
timeout 300 $BENCH_SMALL > gpurun_out/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches.csv $BENCH_SMALL > gpurun_out/ncu1.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'voigt_weights|contract_chol' -s 2 -c 2 -o gpurun_out/prof -f $BENCH_SMALL > gpurun_out/ncu2.log 2>&1
echo "ncu full rc=$?"
tail -3 gpurun_out/ncu2.log
fi
