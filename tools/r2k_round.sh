#!/bin/bash
mkdir -p gpurun_out
bash tools/k1_variants_run.sh 400000 > gpurun_out/r2k_variants.txt 2>&1; cat gpurun_out/r2k_variants.txt
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -x > gpurun_out/r2k_pytest.log 2>&1
echo "pytest exit $?"; tail -4 gpurun_out/r2k_pytest.log
small="--frames 200000 --steps 2 --warmup 1 --no-extras --no-cpu-baseline --e2e-frames 2000 --sustain-s 0"
python bench.py $small > gpurun_out/r2k_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k 'regex:k_site_states' -c 1 -f \
    -o gpurun_out/r2k_k1 python bench.py $small > gpurun_out/r2k_ncu.log 2>&1
echo "ncu k1 $?"
