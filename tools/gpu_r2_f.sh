#!/bin/bash
# round 2, call f: whole GPU suite after the C-ABI guard / MEX / shortcut changes, then the bench line
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f_pytest.log
tail -5 gpurun_out/f_pytest.log
timeout 600 python bench.py > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/f_bench.json; tail -5 gpurun_out/f_bench.err
