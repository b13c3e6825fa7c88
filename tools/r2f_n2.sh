#!/bin/bash
# 2-GPU smoke of the frame-sharded bench (all extras), before the 8-GPU scaling run
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 \
   bench.py --gpus 2 --steps 5 --warmup 3 --phases > gpurun_out/r2f_n2.json 2> gpurun_out/r2f_n2.err
echo "n2 exit $?"; tail -c 2500 gpurun_out/r2f_n2.err; head -c 6000 gpurun_out/r2f_n2.json
