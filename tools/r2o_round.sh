#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q --timeout 600 -k "density or meanfreq or hotpath_optimistic" > gpurun_out/r2o_pytest.log 2>&1
echo "pytest exit $?"; tail -5 gpurun_out/r2o_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
for sh in 20 21 22; do echo "chunk shift $sh"; GD_CHUNK_SHIFT=$sh python tools/step_ab.py 125000 2>&1 | tail -1; done
