#!/bin/bash
# SHARED_DATA plans on config 1 (and a C2 cross-check), the fixed peaks microbenchmark, quantise timing on C3
mkdir -p gpurun_out; L=gpurun_out/c1_blocked.log; : > $L
./tools/microbench/peaks > gpurun_out/peaks.json 2>&1; cat gpurun_out/peaks.json
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/pytest_parity.log 2>&1; tail -5 gpurun_out/pytest_parity.log
echo "== C1 default plan" >> $L; timeout 200 python tools/run_one.py C1 2 0 4 >> $L 2>&1
echo "== C1 TAHOE_SD_BLOCKED=0" >> $L; TAHOE_SD_BLOCKED=0 timeout 200 python tools/run_one.py C1 2 0 4 >> $L 2>&1
for top in 1 3 5 7; do
  echo "== C1 blocked top=$top" >> $L; TAHOE_SD_TOP_LEVELS=$top timeout 200 python tools/run_one.py C1 2 0 4 >> $L 2>&1
done
echo "== C2 1M rows default / unblocked" >> $L
timeout 200 python tools/run_one.py C2 2 1000000 3 >> $L 2>&1
TAHOE_SD_BLOCKED=0 timeout 200 python tools/run_one.py C2 2 1000000 3 >> $L 2>&1
echo "== C1 small batches (1000 rows) default / unblocked" >> $L
timeout 200 python tools/run_one.py C1 2 1000 4 >> $L 2>&1
TAHOE_SD_BLOCKED=0 timeout 200 python tools/run_one.py C1 2 1000 4 >> $L 2>&1
echo "== C3 2M rows quantise + traverse" >> $L
timeout 300 python tools/probe_prof.py C3 2000000 >> $L 2>&1
cat $L
