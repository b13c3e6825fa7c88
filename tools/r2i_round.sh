#!/bin/bash
mkdir -p gpurun_out
small="--frames 200000 --steps 2 --warmup 1 --no-extras --no-cpu-baseline --e2e-frames 2000 --sustain-s 0"
python bench.py $small > gpurun_out/r2i_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k 'regex:k_site_states|k_events_compact|k_jumps_spec' -c 3 -f \
    -o gpurun_out/r2i_k1 python bench.py $small > gpurun_out/r2i_ncu.log 2>&1
echo "ncu k1 $?"
