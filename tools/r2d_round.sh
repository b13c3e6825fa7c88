#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2d_pytest.log 2>&1
echo "pytest exit $?"; tail -12 gpurun_out/r2d_pytest.log
python tools/k0_sweep.py 400000 > gpurun_out/r2d_k0_sweep.txt 2>&1; cat gpurun_out/r2d_k0_sweep.txt
python tools/extras_probe.py 200000 2048 > gpurun_out/r2d_probe.txt 2>&1; echo "probe $?"; cat gpurun_out/r2d_probe.txt
python tools/rdf_time.py > gpurun_out/r2d_rdf_time.txt 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_rdf_hist -c 2 -f -o gpurun_out/r2d_rdf \
    python tools/rdf_time.py > gpurun_out/r2d_ncu_rdf.log 2>&1
echo "ncu rdf $?"; cat gpurun_out/r2d_rdf_time.txt
PROBE_REPS=0 python tools/extras_probe.py 200000 64 > /dev/null 2>&1 && \
PROBE_REPS=0 ncu --set full --clock-control none --import-source on \
    -k "regex:k_density_runs|k_fill_|k_amp_carry|k_unwrap_fused" -c 8 -f \
    -o gpurun_out/r2d_extras python tools/extras_probe.py 200000 64 > gpurun_out/r2d_ncu_extras.log 2>&1
echo "ncu extras $?"
timeout 900 python bench.py > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo "bench $?"; tail -c 1500 gpurun_out/r2d_bench.err
head -c 9000 gpurun_out/r2d_bench.json
