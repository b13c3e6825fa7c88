#!/bin/bash
# round 2c: full GPU parity suite (all failures listed), timings + ncu of the secondary kernels, bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2c_pytest.log 2>&1
echo "pytest exit $?"; tail -25 gpurun_out/r2c_pytest.log
python tools/extras_probe.py 200000 2048 > gpurun_out/r2c_probe.txt 2>&1; echo "probe $?"; cat gpurun_out/r2c_probe.txt
PROBE_REPS=0 python tools/extras_probe.py 200000 2048 > /dev/null 2>&1 && \
PROBE_REPS=0 ncu --set full --clock-control none --import-source on \
    -k "regex:k_density_runs|k_msd|k_fill_|k_occupancy|k_amp_walk|k_unwrap_fused|k_to_soa" -c 24 -f \
    -o gpurun_out/r2c_extras python tools/extras_probe.py 200000 2048 > gpurun_out/r2c_ncu_extras.log 2>&1
echo "ncu extras $?"
python tools/rdf_time.py > gpurun_out/r2c_rdf_time.txt 2>&1; cat gpurun_out/r2c_rdf_time.txt
timeout 900 python bench.py > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; echo "bench $?"; tail -c 1500 gpurun_out/r2c_bench.err
head -c 6000 gpurun_out/r2c_bench.json
