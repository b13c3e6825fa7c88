#!/bin/bash
# K1 variants: band counters through atomics (b1), state changes deferred to a per-32-row flush
# (d1), at 6 and 7 CTAs per SM.  The variants are gd_states.cu with profiles/r2u_k1_defer.patch applied,
# compiled with -DGD_K1_BAND_ATOMIC=1 / -DGD_K1_DEFER_EVENTS=1 / -DGD_K1_MIN_CTAS=7 and linked as in
# tools/k1_variants.sh; timed on the box, then the GPU
# suite under the two candidates, then instruction counts of base and candidate.
mkdir -p gpurun_out
out=gpurun_out/r2u_k1_defer.txt
: > $out
for so in gemdat_b200/csrc/_variants/lib_*.so; do
  n=$(basename $so .so)
  echo -n "$n: " >> $out; GEMDAT_B200_LIB=$PWD/$so timeout 300 python tools/k1_sweep.py 400000 2>&1 | tail -1 >> $out
done
cat $out
for n in b1d1m7 b1d1; do
  GEMDAT_B200_LIB=$PWD/gemdat_b200/csrc/_variants/lib_$n.so timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r2u_pytest_$n.txt 2>&1
  echo "pytest $n exit $?"; tail -2 gpurun_out/r2u_pytest_$n.txt
done
for n in base b1d1m7; do
  GEMDAT_B200_LIB=$PWD/gemdat_b200/csrc/_variants/lib_$n.so timeout 600 ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,sm__warps_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum \
    --clock-control none -k regex:k_site_states -c 2 --csv --log-file gpurun_out/r2u_ncu_$n.csv python tools/k1_sweep.py 200000 > gpurun_out/r2u_ncu_$n.log 2>&1
  echo "ncu $n exit $?"; grep -c k_site_states gpurun_out/r2u_ncu_$n.csv
done
