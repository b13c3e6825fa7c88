#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu -k "displacement or unwrap or shards or tracer or metrics or drift or msd" > gpurun_out/r2zb_pytest.txt 2>&1
echo "k0 tests exit $?"; tail -2 gpurun_out/r2zb_pytest.txt
timeout 600 python tools/extras_probe.py 400000 256 > gpurun_out/r2zb_probe.txt 2>&1; head -4 gpurun_out/r2zb_probe.txt
GD_UNWRAP_CTAS_PER_SM=8 timeout 600 python tools/extras_probe.py 400000 256 2>&1 | head -3
