#!/bin/bash
mkdir -p gpurun_out
bash tools/k1_variants_run.sh 400000 > gpurun_out/r2p_variants.txt 2>&1; cat gpurun_out/r2p_variants.txt
