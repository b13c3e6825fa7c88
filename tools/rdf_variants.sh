#!/bin/bash
# A/B builds of the RDF kernel (-D overrides of the knobs at the top of gd_rdf.cu).
# usage: bash tools/rdf_variants.sh     -> gemdat_b200/csrc/_variants/lib_rdf_<name>.so
# on the box:  for v in base select ctas1 ctas2 ctas3 select_ctas2; do
#                GEMDAT_B200_LIB=gemdat_b200/csrc/_variants/lib_rdf_$v.so python tools/rdf_time.py; done
# and the parity tests with the same variable:  python -m pytest tests -m gpu -k "rdf or radial"
set -euo pipefail
cd "$(dirname "$0")/../gemdat_b200/csrc"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS=(-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 --fmad=false -Xcompiler -fPIC,-ffp-contract=off)
mkdir -p _variants
build() { # name, defines...
  local name=$1; shift
  "$NVCC" "${FLAGS[@]}" "$@" -Xptxas -v -c gd_rdf.cu -o _variants/gd_rdf_$name.o > _variants/rdf_$name.log 2>&1
  objs=$(ls _build/gd_*.o | grep -v gd_rdf.o)
  "$NVCC" -shared -o _variants/lib_rdf_$name.so _variants/gd_rdf_$name.o $objs -lcudart
  echo "rdf_$name: $(grep -A3 'Function properties for .*k_rdf_hist' _variants/rdf_$name.log | grep -o 'Used [0-9]* registers' | head -1)"
}
build base &
build select -DGD_RDF_DIAG_SELECT=1 &
build ctas1 -DGD_RDF_MIN_CTAS=1 &   # explicit 1 frees the allocator: 114 registers instead of 94
build ctas2 -DGD_RDF_MIN_CTAS=2 &
build ctas3 -DGD_RDF_MIN_CTAS=3 &
build select_ctas2 -DGD_RDF_DIAG_SELECT=1 -DGD_RDF_MIN_CTAS=2 &
wait
