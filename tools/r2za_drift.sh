#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu -k "drift or center_of_mass" > gpurun_out/r2za_pytest.txt 2>&1
echo "drift tests exit $?"; tail -2 gpurun_out/r2za_pytest.txt
timeout 600 python tools/extras_probe.py 400000 256 > gpurun_out/r2za_probe.txt 2>&1; cat gpurun_out/r2za_probe.txt | head -30
