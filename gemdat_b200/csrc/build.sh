#!/bin/bash
# Build libgemdat_b200.so in-tree for sm_100a (cross-compiles without a GPU).
set -euo pipefail
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS=(-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17
       --fmad=false -Xcompiler -fPIC,-ffp-contract=off,-Wall -Xptxas -v)
mkdir -p _build
pids=()
for f in gd_*.cu; do
  o=_build/${f%.cu}.o
  stale=0
  for dep in "$f" gd_*.cuh ../../include/gemdat_b200.h build.sh; do [[ $dep -nt $o ]] && stale=1; done
  if [[ ! -f $o || $stale == 1 ]]; then
    ( "$NVCC" "${FLAGS[@]}" -c "$f" -o "$o" > "_build/${f%.cu}.log" 2>&1 || { cat "_build/${f%.cu}.log"; exit 1; } ) &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [[ -n "$p" ]] && wait "$p"; done
# link beside the target, then rename: a reader (or a gpurun snapshot) never sees a partial file
"$NVCC" -shared -o _build/libgemdat_b200.so.tmp _build/gd_*.o -lcudart
mv -f _build/libgemdat_b200.so.tmp libgemdat_b200.so
echo "built $(pwd)/libgemdat_b200.so"
