#!/bin/bash
# 1/2/4/8-GPU strong-scaling pass of bench.py on one box (under `gpurun --gpus 8`).
# usage: bash tools/scale_run.sh <tag> [extra bench args]
tag=${1:-scale}; shift
out=gpurun_out; mkdir -p $out
args="--no-extras --no-cpu-baseline --phases --steps 5 $*"
for n in 1 2 4 8; do
  if [[ $n == 1 ]]; then
    timeout 300 python bench.py --gpus 1 $args > $out/${tag}_n$n.json 2> $out/${tag}_n$n.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
      --master-port $((29500 + n)) bench.py --gpus $n $args > $out/${tag}_n$n.json 2> $out/${tag}_n$n.err
  fi
  echo "n=$n exit $?"; tail -c 300 $out/${tag}_n$n.json | head -c 300; echo
done
