#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2l_pytest.log 2>&1
echo "pytest exit $?"; tail -8 gpurun_out/r2l_pytest.log
timeout 900 python bench.py --no-extras --no-cpu-baseline --e2e-frames 2000 --sustain-s 0 > gpurun_out/r2l_bench.json 2> gpurun_out/r2l_bench.err; echo "bench $?"; tail -c 600 gpurun_out/r2l_bench.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2l_bench.json') if l.startswith('{')][-1])
print('ms', d['ms_per_step'], 'first_call', d['first_call'])
PY
GD_FIRST_CALL_BOUND=0 timeout 900 python bench.py --no-extras --no-cpu-baseline --e2e-frames 2000 --sustain-s 0 2>/dev/null | python -c "
import sys,json
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('checked first call', d['first_call'])"
