#!/bin/bash
# on the GPU box: times K1 for every variant library built by k1_variants.sh
for so in gemdat_b200/csrc/_variants/lib_*.so; do
  n=$(basename $so .so)
  echo -n "$n staged: "; GEMDAT_B200_LIB=$PWD/$so timeout 120 python scratch/k1_sweep.py 400000 2>&1 | tail -1
done
echo -n "base reread: "; GD_EVENTS_STAGED=0 GEMDAT_B200_LIB=$PWD/gemdat_b200/csrc/_variants/lib_base.so timeout 120 python scratch/k1_sweep.py 400000 2>&1 | tail -1
