#!/bin/bash
# K1 ring slots of GD_K1_TILE rows (slot bookkeeping once per tile): variants built with
# -DGD_K1_TILE / -DGD_K1_STAGES / -DGD_K1_MIN_CTAS into gemdat_b200/csrc/_variants (see tools/k1_variants.sh)
mkdir -p gpurun_out
out=gpurun_out/r2y_k1_tile.txt; : > $out
for so in gemdat_b200/csrc/_variants/lib_*.so; do
  n=$(basename $so .so)
  echo -n "$n: " >> $out; GEMDAT_B200_LIB=$PWD/$so timeout 300 python tools/k1_sweep.py 400000 2>&1 | tail -1 >> $out
done
cat $out
best=$(python - <<'PY'
import json,re
best=None
for l in open('gpurun_out/r2y_k1_tile.txt'):
    n,_,j=l.partition(': ')
    try: ms=json.loads(j)['k1_ms']
    except Exception: continue
    if n!='lib_base' and (best is None or ms<best[1]): best=(n,ms)
print(best[0])
PY
)
echo "best variant: $best"
GEMDAT_B200_LIB=$PWD/gemdat_b200/csrc/_variants/$best.so timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r2y_pytest_$best.txt 2>&1
echo "pytest $best exit $?"; tail -2 gpurun_out/r2y_pytest_$best.txt
for n in lib_base $best; do
  GEMDAT_B200_LIB=$PWD/gemdat_b200/csrc/_variants/$n.so timeout 600 ncu --metrics smsp__inst_executed.sum,gpu__time_duration.sum,sm__warps_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum \
    --clock-control none -k regex:k_site_states -c 2 --csv --log-file gpurun_out/r2y_ncu_$n.csv python tools/k1_sweep.py 200000 > gpurun_out/r2y_ncu_$n.log 2>&1
  echo "ncu $n exit $?"; grep k_site_states gpurun_out/r2y_ncu_$n.csv | awk -F'","' '{print $(NF-2), $(NF-1), $NF}' | head -5
done
