#!/bin/bash
# round 2, call A: suite, C3 plan sweep with one-copy TMA tiles, new bench line
mkdir -p gpurun_out; L=gpurun_out/r2a_plans.log; : > $L
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2a_pytest.log
python -m tahoe_b200.peaks > gpurun_out/r2a_peaks.json 2>&1; cat gpurun_out/r2a_peaks.json
for plan in 8,1,6 8,0,5 8,0,4 4,0,6 16,0,3 16,1,5; do
  TAHOE_SMEM_PLAN=$plan timeout 300 python tools/probe_prof.py C3 2000000 >> $L 2>&1
done
for plan in 8,1,6 8,0,5; do
  TAHOE_SMEM_PLAN=$plan timeout 300 python tools/probe_prof.py C3 10000000 >> $L 2>&1
done
for bb in 48000000 96000000; do
  TAHOE_ROW_BATCH_BYTES=$bb TAHOE_SMEM_PLAN=8,0,5 timeout 300 python tools/probe_prof.py C3 10000000 >> $L 2>&1
done
cut -c47-300 $L
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"; tail -c 600 gpurun_out/r2a_bench.err; head -c 1500 gpurun_out/r2a_bench.json
