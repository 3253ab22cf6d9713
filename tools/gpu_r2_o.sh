#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_slices.py tests/test_gpu_parity.py tests/test_gpu_parity_r2.py -m gpu -x -q > gpurun_out/o_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/o_pytest.log
for i in 1 2; do
timeout 600 python bench.py --workload config5 --quasars 16 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/o_cfg5.json 2> gpurun_out/o_cfg5.err; echo "cfg5 rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/o_cfg5.json')); print('config5', d['value'], d['kernel_ms_per_step'])"
done
bash tools/gpu_ab.sh "-DDLA_K2C_ROWS=16 -DDLA_K2C_MINB=5"
