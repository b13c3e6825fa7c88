#!/bin/bash
# Builds A/B variants of libgemdat_b200.so that differ only in K1 tuning knobs.
# usage: bash scratch/k1_variants.sh   -> gemdat_b200/csrc/_variants/lib_<name>.so
set -euo pipefail
cd "$(dirname "$0")/../gemdat_b200/csrc"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS=(-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --fmad=false -Xcompiler -fPIC,-ffp-contract=off)
mkdir -p _variants
build() { # name, defines...
  local name=$1; shift
  "$NVCC" "${FLAGS[@]}" "$@" -Xptxas -v -c gd_states.cu -o _variants/gd_states_$name.o > _variants/$name.log 2>&1
  objs=$(ls _build/gd_*.o | grep -v gd_states.o)
  "$NVCC" -shared -o _variants/lib_$name.so _variants/gd_states_$name.o $objs -lcudart
  echo "$name: $(grep -A1 'k_site_statesILb1ELb1ELb1ELb1' _variants/$name.log | grep -o 'Used [0-9]* registers' | head -1) $(grep -B1 'Used' _variants/$name.log | grep -o '[0-9]* bytes spill stores' | sort -u | tr '\n' ' ')"
}
build base &
build nofilter -DGD_K1_IMAGE_FILTER=0 &
build w4 -DGD_K1_WARPS=4 -DGD_K1_MIN_CTAS=6 &
wait
