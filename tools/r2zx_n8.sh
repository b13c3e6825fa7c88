#!/bin/bash
# final tree on 8 GPUs: the N=8 line (no extras) -- also the check that the multi-rank path still runs
mkdir -p gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29508 \
   bench.py --gpus 8 --steps 5 --warmup 3 --phases --no-extras --no-c2 --no-cpu-baseline > gpurun_out/r2zx_n8.json 2> gpurun_out/r2zx_n8.err
echo "n8 exit $?"
python - <<'PY'
import json
for l in open('gpurun_out/r2zx_n8.json'):
    if l.startswith('{'):
        d = json.loads(l); print('ms', d['ms_per_step'], 'value', d['value'], 'first', d['first_call']['ms'], 'e2e', d['e2e']['value'], d['e2e']['ms_per_step'], d.get('phases_ms'))
PY
tail -3 gpurun_out/r2zx_n8.err
