#!/bin/bash
# after the one-pass SiteGrid: the tests that build grids through every route, then the bench line
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu -k "pipeline or pipelined or golden or hotpath or c3_two_label" > gpurun_out/r2zz_pytest.txt 2>&1
echo "tests exit $?"; tail -2 gpurun_out/r2zz_pytest.txt
timeout 600 python bench.py --no-c2 --no-extras > gpurun_out/r2zz_bench_c3.json 2> gpurun_out/r2zz_bench_c3.err
echo "bench exit $?"
python - <<'PY'
import json
for l in open('gpurun_out/r2zz_bench_c3.json'):
    if l.startswith('{'):
        d = json.loads(l); print('ms_per_step', d['ms_per_step'], 'first', d['first_call']['ms'], 'e2e', d['e2e']['value'], d['e2e']['breakdown_ms'])
PY
