#!/bin/bash
# paired-warp variant of k_smem_forest_deep: sparse parity tests, then plan x depth sweep on config 4
mkdir -p gpurun_out; L=gpurun_out/pair.log; : > $L
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "sparse" > gpurun_out/pytest_pair.log 2>&1; tail -3 gpurun_out/pytest_pair.log
for plan in "16,1,6" "16,1,4" "8,1,6"; do
  for d in 6 7 8; do
    TAHOE_SMEM_PLAN=$plan TAHOE_QTOP_DEPTH=$d timeout 300 python tools/probe_qtop.py 1000000 >> $L 2>&1
  done
done
cut -c1-40,60-70,100-400 $L
