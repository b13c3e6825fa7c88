#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -x > gpurun_out/r2j_pytest.log 2>&1
echo "pytest exit $?"; tail -4 gpurun_out/r2j_pytest.log
GD_COMPACT_CTAS_PER_SM=100000 python tools/step_ab.py 1000000 2>&1 | head -1 > gpurun_out/r2j_step.txt; cat gpurun_out/r2j_step.txt
small="--frames 200000 --steps 2 --warmup 1 --no-extras --no-cpu-baseline --e2e-frames 2000 --sustain-s 0"
python bench.py $small > gpurun_out/r2j_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k 'regex:k_site_states' -c 1 -f \
    -o gpurun_out/r2j_k1 python bench.py $small > gpurun_out/r2j_ncu.log 2>&1
echo "ncu k1 $?"
