#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 600 -k "attempt or displacement or collective_through" > gpurun_out/r2n_pytest.log 2>&1
echo "pytest exit $?"; tail -15 gpurun_out/r2n_pytest.log
python tools/extras_probe.py 200000 64 2>&1 | grep -E "N2|speed" 
