#!/bin/bash
# Builds A/B variants of libgemdat_b200.so that differ only in tuning knobs (-D overrides) or in
# the source of one translation unit.
# usage: bash tools/k1_variants.sh   -> gemdat_b200/csrc/_variants/lib_<name>.so
set -euo pipefail
cd "$(dirname "$0")/../gemdat_b200/csrc"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS=(-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --fmad=false -Xcompiler -fPIC,-ffp-contract=off -I.)
mkdir -p _variants
rm -f _variants/lib_*.so
build() { # name, source file, source stem it replaces (gd_states | gd_events | gd_jumps), defines...
  local name=$1 src=$2 stem=$3; shift 3
  "$NVCC" "${FLAGS[@]}" "$@" -Xptxas -v -c $src -o _variants/${stem}_$name.o > _variants/$name.log 2>&1
  objs=$(ls _build/gd_*.o | grep -v "/$stem.o")
  "$NVCC" -shared -o _variants/lib_$name.so _variants/${stem}_$name.o $objs -lcudart
  echo "$name built: $(grep -A1 'k_site_statesILb1ELb1ELb1ELb1' _variants/$name.log | grep -o 'Used [0-9]* registers' | head -1)"
}
build base gd_states.cu gd_states &
build min7 gd_states.cu gd_states -DGD_K1_MIN_CTAS=7 &
build min8 gd_states.cu gd_states -DGD_K1_MIN_CTAS=8 &
build min8w2 gd_states.cu gd_states -DGD_K1_MIN_CTAS=16 -DGD_K1_WARPS=2 &
wait
