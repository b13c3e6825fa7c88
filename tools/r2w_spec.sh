#!/bin/bash
# K3 speculative pass: lane-per-segment streaming (GD_K3_SPEC_TILED=0) against the cp.async tile kernel (=1)
mkdir -p gpurun_out
out=gpurun_out/r2w_spec_ab.txt; : > $out
for w in 0 1 0 1; do GD_K3_SPEC_TILED=$w timeout 300 python tools/step_ab.py 1000000 2>&1 | tail -1 >> $out; done
for w in 0 1; do GD_K3_SPEC_TILED=$w timeout 300 python tools/step_ab.py 125000 2>&1 | tail -1 >> $out; done
cat $out
GD_K3_SPEC_TILED=1 timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r2w_pytest_tiled.txt 2>&1
echo "pytest exit $?"; tail -2 gpurun_out/r2w_pytest_tiled.txt
GD_K3_SPEC_TILED=1 timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum \
  --clock-control none -k regex:k_jumps_spec -c 2 --csv --log-file gpurun_out/r2w_ncu_tiled.csv python tools/step_ab.py 1000000 > gpurun_out/r2w_ncu.log 2>&1
echo "ncu exit $?"; grep k_jumps_spec gpurun_out/r2w_ncu_tiled.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}'
