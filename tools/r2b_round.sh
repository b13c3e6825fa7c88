#!/bin/bash
# round 2b: full GPU parity suite, then baseline ncu captures of the secondary kernels
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/r2b_pytest.log 2>&1
echo "pytest exit $?"; tail -15 gpurun_out/r2b_pytest.log
bash tools/r2a_probe.sh
