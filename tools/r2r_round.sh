#!/bin/bash
mkdir -p gpurun_out
for T in 125000 250000 500000 1000000; do for tune in 0 1; do echo -n "frames $T tune $tune: "; GD_CHUNK_TUNE=$tune python tools/step_ab.py $T 2>&1 | tail -1; done; done > gpurun_out/r2r_chunk_tune.txt 2>&1
cat gpurun_out/r2r_chunk_tune.txt
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2r_pytest.log 2>&1
echo "pytest exit $?"; tail -5 gpurun_out/r2r_pytest.log
