#!/bin/bash
# last check of the round: smoke(), the full GPU suite, the default bench line
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2zy_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/r2zy_smoke.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2zy_pytest_gpu.txt 2>&1
echo "pytest gpu exit $?"; tail -2 gpurun_out/r2zy_pytest_gpu.txt
timeout 900 python bench.py > gpurun_out/r2zy_bench_c3.json 2> gpurun_out/r2zy_bench_c3.err
echo "bench exit $?"
python - <<'PY'
import json
for l in open('gpurun_out/r2zy_bench_c3.json'):
    if l.startswith('{'):
        d = json.loads(l); print('ms_per_step', d['ms_per_step'], 'first', d['first_call']['ms'], 'e2e', d['e2e']['value'], 'roofline', d['roofline']['frac'])
PY
