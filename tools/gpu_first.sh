#!/bin/bash
# first GPU session: FP64 probes, gpu tests, a short bench
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt; which octave mkoctfile >> gpurun_out/nproc.txt 2>&1
timeout 120 ./tools/probe_fp64 > gpurun_out/probe_fp64.jsonl 2>&1; echo "probe rc=$?"
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -40 > gpurun_out/pytest_gpu.log; echo "pytest rc=${PIPESTATUS[0]}"
tail -30 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 3 --warmup 3 --quasars 64 > gpurun_out/bench_first.json 2> gpurun_out/bench_first.err; echo "bench rc=$?"
cat gpurun_out/bench_first.json; tail -5 gpurun_out/bench_first.err
cat gpurun_out/probe_fp64.jsonl
