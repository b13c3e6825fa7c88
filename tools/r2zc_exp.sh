#!/bin/bash
mkdir -p gpurun_out
out=gpurun_out/r2zc_k1_flags.txt; : > $out
for cfg in "" "GD_EXP_NO_BAND=1" "GD_EXP_NO_VIS=1" "GD_EXP_NO_BAND=1 GD_EXP_NO_VIS=1" ""; do
  echo -n "[$cfg] " >> $out; env $cfg timeout 300 python tools/k1_sweep.py 400000 2>&1 | tail -1 >> $out
done
cat $out
