#!/bin/bash
mkdir -p gpurun_out
python tools/extras_probe.py 200000 256 > gpurun_out/r2m_probe.txt 2>&1; echo "probe $?"; cat gpurun_out/r2m_probe.txt
