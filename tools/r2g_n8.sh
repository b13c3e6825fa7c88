#!/bin/bash
# 8-GPU box: N=8 and N=4 bench lines (all stages frame-sharded), then BASELINE config 4 (all lags) at N=8
mkdir -p gpurun_out
run() { n=$1; tag=$2; shift 2
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
     bench.py --gpus $n "$@" > gpurun_out/${tag}.json 2> gpurun_out/${tag}.err
  echo "$tag exit $?"; tail -c 600 gpurun_out/${tag}.err | tail -3; grep '^{' gpurun_out/${tag}.json | head -c 700; echo; }
run 8 r2g_n8 --steps 5 --warmup 3 --phases
run 4 r2g_n4 --steps 5 --warmup 3 --phases
run 8 r2g_n8_alllags --steps 2 --warmup 3 --msd-lags 0 --no-cpu-baseline --no-c2 --e2e-frames 2000 --sustain-s 0
