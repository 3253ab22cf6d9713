#!/bin/bash
# round 2, call A: all GPU tests (not -x: every failure is wanted), the full-size parity subsamples, the default bench
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -s 2>&1 | tail -120 > gpurun_out/pytest_gpu.log; echo "pytest rc=${PIPESTATUS[0]}"
tail -40 gpurun_out/pytest_gpu.log
timeout 900 python tools/parity_full.py > gpurun_out/parity_full.log 2>&1; echo "parity_full rc=$?"
tail -8 gpurun_out/parity_full.log
timeout 600 python bench.py > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"
cat gpurun_out/bench_default.json
