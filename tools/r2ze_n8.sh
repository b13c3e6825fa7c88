#!/bin/bash
# 8-GPU box: the N=8 and N=4 lines after the K2 compaction change (no extras: the stage table of
# DESIGN section 5 keeps its earlier lines), then N=2
mkdir -p gpurun_out
run() { n=$1; tag=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
     bench.py --gpus $n "$@" > gpurun_out/${tag}.json 2> gpurun_out/${tag}.err
  echo "$tag exit $?"; python - "$tag" <<'PY'
import json, sys
for l in open(f'gpurun_out/{sys.argv[1]}.json'):
    if l.startswith('{'):
        d = json.loads(l); print(sys.argv[1], 'ms', d['ms_per_step'], 'value', d['value'], 'first', d['first_call']['ms'], 'e2e', d['e2e']['value'], d.get('phases_ms'))
PY
}
run 8 r2ze_n8 --steps 5 --warmup 3 --phases --no-extras --no-c2 --no-cpu-baseline
run 4 r2ze_n4 --steps 5 --warmup 3 --phases --no-extras --no-c2 --no-cpu-baseline
run 2 r2ze_n2 --steps 5 --warmup 3 --phases --no-extras --no-c2 --no-cpu-baseline
