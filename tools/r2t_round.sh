#!/bin/bash
# round 2, final tree: the allocation-free entry point alone in a fresh process, the full GPU
# suite, the default bench line, the reference arm, and the ncu launch list of the bench command
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r2t_noalloc_alone.txt 2>&1
echo "noalloc alone exit $?"; tail -3 gpurun_out/r2t_noalloc_alone.txt
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2t_pytest_gpu.txt 2>&1
echo "pytest gpu exit $?"; tail -5 gpurun_out/r2t_pytest_gpu.txt
timeout 900 python bench.py > gpurun_out/r2t_bench_c3.json 2> gpurun_out/r2t_bench_c3.err
rc=$?; echo "bench exit $rc"; head -c 600 gpurun_out/r2t_bench_c3.json; echo
python - <<'PY'
import json
for l in open('gpurun_out/r2t_bench_c3.json'):
    if l.startswith('{'):
        d = json.loads(l); print('ms_per_step', d['ms_per_step'], 'e2e', d['e2e']['value'], d['e2e']['pageable'])
PY
timeout 600 python bench.py --impl reference > gpurun_out/r2t_bench_reference.json 2> gpurun_out/r2t_bench_reference.err
echo "reference exit $?"; head -c 500 gpurun_out/r2t_bench_reference.json; echo
if [[ $rc == 0 ]]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv \
    --log-file gpurun_out/r2t_launches.csv python bench.py --steps 2 --warmup 3 --no-c2 --no-cpu-baseline --no-extras \
    > gpurun_out/r2t_ncu.log 2>&1
  echo "ncu exit $?"; wc -l gpurun_out/r2t_launches.csv
fi
