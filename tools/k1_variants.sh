#!/bin/bash
# Builds A/B variants of libgemdat_b200.so that differ only in tuning knobs (-D overrides).
# usage: bash tools/k1_variants.sh   -> gemdat_b200/csrc/_variants/lib_<name>.so
set -euo pipefail
cd "$(dirname "$0")/../gemdat_b200/csrc"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS=(-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --fmad=false -Xcompiler -fPIC,-ffp-contract=off)
mkdir -p _variants
build() { # name, source stem (gd_states | gd_events | gd_jumps), defines...
  local name=$1 stem=$2; shift 2
  "$NVCC" "${FLAGS[@]}" "$@" -Xptxas -v -c $stem.cu -o _variants/${stem}_$name.o > _variants/$name.log 2>&1
  objs=$(ls _build/gd_*.o | grep -v $stem.o)
  "$NVCC" -shared -o _variants/lib_$name.so _variants/${stem}_$name.o $objs -lcudart
  echo "$name built"
}
build base gd_states &
build novis gd_states -DGD_K1_PROBE_NO_VIS=1 &
wait
