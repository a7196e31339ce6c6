#!/bin/bash
# round 2, call F: 16 row teams on config 1; reorder / simhash ablation; same-box reference kernels at large batches
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2f_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2f_pytest.log
timeout 200 python tools/probe_c1.py 2>&1 | tail -1
A=gpurun_out/r2f_reorder.log; : > $A
for shape in "500 10 18 200000" "1000 12 28 100000" "100 8 3072 50000"; do
  timeout 200 python tools/probe_reorder.py $shape >> $A 2>&1
  TAHOE_NO_REORDER=1 timeout 200 python tools/probe_reorder.py $shape >> $A 2>&1
  TAHOE_NO_SIMHASH=1 timeout 200 python tools/probe_reorder.py $shape >> $A 2>&1
  TAHOE_NO_REORDER=1 TAHOE_NO_SIMHASH=1 timeout 200 python tools/probe_reorder.py $shape >> $A 2>&1
done
cat $A
bash tools/gpu_r2/ref_compare_big.sh
