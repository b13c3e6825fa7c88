#!/bin/bash
# K2 compaction: 16 lanes per segment (GD_COMPACT_WARP=0) against the warp-cooperative kernel (=1)
mkdir -p gpurun_out
out=gpurun_out/r2v_compact_ab.txt; : > $out
for w in 0 1 0 1; do GD_COMPACT_WARP=$w timeout 300 python tools/step_ab.py 1000000 2>&1 | tail -1 >> $out; done
for w in 0 1; do GD_COMPACT_WARP=$w timeout 300 python tools/step_ab.py 125000 2>&1 | tail -1 >> $out; done
cat $out
GD_COMPACT_WARP=1 timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r2v_pytest_warp.txt 2>&1
echo "pytest exit $?"; tail -2 gpurun_out/r2v_pytest_warp.txt
