#!/bin/bash
# baseline ncu captures of the secondary kernels (before tuning them)
mkdir -p gpurun_out
python tools/extras_probe.py 200000 2048 > gpurun_out/r2a_probe.txt 2>&1 && \
ncu --set full --clock-control none --import-source on -k "regex:k_density|k_unwrap|k_msd|k_to_soa|k_fill_" -c 12 -f \
    -o gpurun_out/r2a_extras python tools/extras_probe.py 200000 2048 > gpurun_out/r2a_ncu_extras.log 2>&1
echo "extras $?"
python tools/rdf_time.py > gpurun_out/r2a_rdf_time.txt 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_rdf_hist -c 4 -f -o gpurun_out/r2a_rdf \
    python tools/rdf_time.py > gpurun_out/r2a_ncu_rdf.log 2>&1
echo "rdf $?"
cat gpurun_out/r2a_probe.txt gpurun_out/r2a_rdf_time.txt
