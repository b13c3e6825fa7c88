#!/bin/bash
# round 2, after the K2 compaction / K0 dist / drift work: full suite, both bench arms, one full ncu
# capture of K1 + the new compaction kernel + the drift kernels
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2zm_noalloc_alone.txt 2>&1
echo "noalloc alone exit $?"; tail -1 gpurun_out/r2zm_noalloc_alone.txt
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2zm_pytest_gpu.txt 2>&1
echo "pytest gpu exit $?"; tail -3 gpurun_out/r2zm_pytest_gpu.txt
timeout 900 python bench.py > gpurun_out/r2zm_bench_c3.json 2> gpurun_out/r2zm_bench_c3.err
rc=$?; echo "bench exit $rc"; tail -2 gpurun_out/r2zm_bench_c3.err
python - <<'PY'
import json
for l in open('gpurun_out/r2zm_bench_c3.json'):
    if l.startswith('{'):
        d = json.loads(l); print('ms_per_step', d['ms_per_step'], 'value', d['value'], 'first', d['first_call']['ms'], 'e2e', d['e2e']['value'], 'roofline', d['roofline']['frac'], 'launches', d['gpu_launches'])
        print({k: round(v.get('frac_of_hbm', v.get('frac_of_fp64_peak', 0)) or 0, 3) for k, v in d['extras'].items() if isinstance(v, dict) and ('frac_of_hbm' in v or 'frac_of_fp64_peak' in v)})
PY
timeout 600 python bench.py --impl reference > gpurun_out/r2zm_bench_reference.json 2> gpurun_out/r2zm_bench_reference.err
echo "reference exit $?"; head -c 300 gpurun_out/r2zm_bench_reference.json; echo
if [[ $rc == 0 ]]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r2zm_launches.csv \
      python bench.py --steps 2 --warmup 3 --no-c2 --no-cpu-baseline --no-extras > gpurun_out/r2zm_ncu.log 2>&1
  echo "ncu launch list exit $?"; wc -l gpurun_out/r2zm_launches.csv
fi
