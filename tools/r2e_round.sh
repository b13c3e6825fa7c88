#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 600 -k "rdf or radial or density or collective" > gpurun_out/r2e_pytest.log 2>&1
echo "pytest exit $?"; tail -4 gpurun_out/r2e_pytest.log
python tools/rdf_time.py > gpurun_out/r2e_rdf_time.txt 2>&1; cat gpurun_out/r2e_rdf_time.txt
python tools/step_ab.py 1000000 > gpurun_out/r2e_step_ab.txt 2>&1; cat gpurun_out/r2e_step_ab.txt
