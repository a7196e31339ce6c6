#!/bin/bash
# round 2, call D: compile-time parameter-table levels; full suite; C1 team/buffer plans; bench line
mkdir -p gpurun_out; L=gpurun_out/r2d_plans.log; : > $L
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2d_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2d_pytest.log
for tl in 0 2; do
  for plan in 8,1,6 12,1,5 12,0,4 16,1,5 16,0,3; do
    TAHOE_TOP_LEVELS=$tl TAHOE_SMEM_PLAN=$plan timeout 300 python tools/probe_prof.py C3 2000000 >> $L 2>&1
  done
  TAHOE_TOP_LEVELS=$tl timeout 300 python tools/probe_prof.py C2 >> $L 2>&1
  for plan in 16,1,5 16,1,4 16,1,6 12,1,5; do
    TAHOE_TOP_LEVELS=$tl TAHOE_SMEM_PLAN=$plan timeout 300 python tools/probe_prof.py C2 >> $L 2>&1
  done
done
cut -c1-230 $L | sed "s/checksum.*//"
C=gpurun_out/r2d_c1.log; : > $C
timeout 200 python tools/probe_c1.py >> $C 2>&1
TAHOE_SD_TEAM=2 timeout 200 python tools/probe_c1.py >> $C 2>&1
TAHOE_SD_TEAM=2 TAHOE_SD_BUFS=1 timeout 200 python tools/probe_c1.py >> $C 2>&1
TAHOE_SD_TEAM=1 TAHOE_SD_BUFS=1 timeout 200 python tools/probe_c1.py >> $C 2>&1
TAHOE_SD_TEAM=2 TAHOE_SD_BUFS=1 TAHOE_SD_TOP_LEVELS=3 timeout 200 python tools/probe_c1.py >> $C 2>&1
TAHOE_SD_TEAM=2 TAHOE_SD_BUFS=1 TAHOE_SD_TOP_LEVELS=7 timeout 200 python tools/probe_c1.py >> $C 2>&1
TAHOE_SD_TEAM=4 TAHOE_SD_BUFS=1 timeout 200 python tools/probe_c1.py >> $C 2>&1
cat $C
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo "bench rc=$?"; tail -c 400 gpurun_out/r2d_bench.err; head -c 900 gpurun_out/r2d_bench.json
