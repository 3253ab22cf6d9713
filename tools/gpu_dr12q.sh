#!/bin/bash
# BASELINE config 2 at full size: 162 861 quasars x 10 000 hypotheses on N GPUs (tools/run_dr12q.py)
mkdir -p gpurun_out
N=${1:-8}
if [ $N = 1 ]; then
  timeout 1200 python tools/run_dr12q.py --repeat $DR12Q_ARGS > gpurun_out/dr12q_n$N.json 2> gpurun_out/dr12q_n$N.err
else
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 \
    tools/run_dr12q.py --repeat $DR12Q_ARGS > gpurun_out/dr12q_n$N.json 2> gpurun_out/dr12q_n$N.err
fi
echo "rc=$?"; grep '^{' gpurun_out/dr12q_n$N.json | cut -c1-1500; tail -3 gpurun_out/dr12q_n$N.err
