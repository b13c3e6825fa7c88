#!/bin/bash
# new host-side consumers (rates, drift, centre of mass), full GPU suite, bench extras with the drift lines
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu -k "drift or center_of_mass or rates" > gpurun_out/r2z_pytest_new.txt 2>&1
echo "new tests exit $?"; tail -5 gpurun_out/r2z_pytest_new.txt
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/r2z_pytest_gpu.txt 2>&1
echo "pytest gpu exit $?"; tail -3 gpurun_out/r2z_pytest_gpu.txt
timeout 900 python bench.py --no-c2 --no-cpu-baseline > gpurun_out/r2z_bench_c3.json 2> gpurun_out/r2z_bench_c3.err
echo "bench exit $?"; tail -3 gpurun_out/r2z_bench_c3.err
python - <<'PY'
import json
for l in open('gpurun_out/r2z_bench_c3.json'):
    if l.startswith('{'):
        d = json.loads(l); print('ms_per_step', d['ms_per_step'], 'first', d['first_call']['ms'])
        for k in ('unwrap', 'drift_correction', 'center_of_mass'):
            print(k, {a: b for a, b in d['extras'][k].items() if a in ('ms', 'frac_of_hbm', 'algorithmic_GBps', 'in_unit_cell')})
PY
