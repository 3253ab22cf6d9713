#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_slices.py tests/test_gpu_parity_r2.py -m gpu -q -s 2>&1 | tail -60 > gpurun_out/pytest_slices.log; echo "pytest rc=${PIPESTATUS[0]}"
tail -60 gpurun_out/pytest_slices.log
