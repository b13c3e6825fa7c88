#!/bin/bash
# K0 resident CTAs per SM for the 8-byte outputs (dist / speed), and the PCIe duplex check behind e2e
mkdir -p gpurun_out
out=gpurun_out/r2zf_k0_persm.txt; : > $out
for n in 4 5 6 7 8 10; do echo "== GD_UNWRAP_CTAS_PER_SM=$n" >> $out; GD_UNWRAP_CTAS_PER_SM=$n timeout 300 python tools/extras_probe.py 400000 16 2>&1 | head -3 >> $out; done
cat $out
timeout 120 python tools/pcie_duplex.py > gpurun_out/r2zf_pcie.txt 2>&1; cat gpurun_out/r2zf_pcie.txt
