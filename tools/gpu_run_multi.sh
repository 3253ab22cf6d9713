#!/bin/bash
# 2-GPU run of the bench (torchrun, one rank per GPU) + the reference arm
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi_multi.txt
N=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 \
   bench.py --gpus $N --steps 3 --warmup 3 --quasars 64 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench n=$N rc=$?"
cat gpurun_out/bench_n$N.json; tail -3 gpurun_out/bench_n$N.err
timeout 600 python bench.py --impl reference --gpus 1 --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cat gpurun_out/bench_ref.json | cut -c1-600
