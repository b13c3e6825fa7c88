#!/bin/bash
# 8-GPU box: final N=8 / N=4 / N=2 bench lines of the round (all stages frame-sharded)
mkdir -p gpurun_out
run() { n=$1; tag=$2; shift 2
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
     bench.py --gpus $n "$@" > gpurun_out/${tag}.json 2> gpurun_out/${tag}.err
  echo "$tag exit $?"; grep '^{' gpurun_out/${tag}.json | head -c 400; echo; }
run 8 r2s_n8 --steps 5 --warmup 3 --phases
run 4 r2s_n4 --steps 5 --warmup 3 --phases
run 2 r2s_n2 --steps 5 --warmup 3 --phases
