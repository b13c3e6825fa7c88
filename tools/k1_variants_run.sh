#!/bin/bash
# on the GPU box: times K1 and the K2/K3 phases for every variant library built by k1_variants.sh
frames=${1:-400000}
for so in gemdat_b200/csrc/_variants/lib_*.so; do
  n=$(basename $so .so)
  echo -n "$n: "; GEMDAT_B200_LIB=$PWD/$so timeout 120 python tools/k1_sweep.py $frames 2>&1 | tail -1
done
