#!/bin/bash
# K3 segment length (rows per speculative segment): 128 (library) against 64 / 32 (variant libraries
# built with -DGD_K3_SEG), each with the streaming and the tiled speculative kernel
mkdir -p gpurun_out
out=gpurun_out/r2x_kseg.txt; : > $out
run() { echo -n "$1 tiled=$2 frames=$3: " >> $out
  if [[ -n "$4" ]]; then export GEMDAT_B200_LIB=$4; else unset GEMDAT_B200_LIB; fi
  GD_K3_SPEC_TILED=$2 timeout 300 python tools/step_ab.py $3 2>&1 | tail -1 >> $out; unset GEMDAT_B200_LIB; }
for fr in 1000000 125000; do
  for t in 0; do
    run seg128 $t $fr ""
    run seg64 $t $fr $PWD/gemdat_b200/csrc/_variants/lib_seg64.so
    run seg32 $t $fr $PWD/gemdat_b200/csrc/_variants/lib_seg32.so
  done
done
cat $out
for sg in 64 32; do
GEMDAT_B200_LIB=$PWD/gemdat_b200/csrc/_variants/lib_seg$sg.so timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r2x_pytest_seg$sg.txt 2>&1
echo "pytest seg$sg exit $?"; tail -2 gpurun_out/r2x_pytest_seg$sg.txt
done
