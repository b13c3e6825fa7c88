#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2h_pytest.log 2>&1
echo "pytest exit $?"; tail -6 gpurun_out/r2h_pytest.log
python tools/k0_sweep.py 400000 > gpurun_out/r2h_k0_sweep.txt 2>&1; cat gpurun_out/r2h_k0_sweep.txt
python tools/step_ab.py 125000 > gpurun_out/r2h_step_125k.txt 2>&1; tail -1 gpurun_out/r2h_step_125k.txt
timeout 900 python bench.py --no-c2 --no-cpu-baseline > gpurun_out/r2h_bench.json 2> gpurun_out/r2h_bench.err; echo "bench $?"; tail -c 800 gpurun_out/r2h_bench.err
grep '^{' gpurun_out/r2h_bench.json | head -c 1200
