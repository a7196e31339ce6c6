#!/bin/bash
# round 2, call U: three table levels (TAHOE_TOP_LEVELS=3) against two on the dense shared-memory-forest kernel
mkdir -p gpurun_out
for tl in 2 3; do
  for plan in auto 16,1,5 16,0,6; do
    TAHOE_TOP_LEVELS=$tl TAHOE_SMEM_PLAN=$plan python tools/probe_prof.py C2 2>&1 | tail -1
  done
  for plan in auto 16,1,5 12,1,5; do
    TAHOE_TOP_LEVELS=$tl TAHOE_SMEM_PLAN=$plan python tools/probe_prof.py C3 2000000 2>&1 | tail -1
  done
done | tee gpurun_out/r2u_plans.log
TAHOE_TOP_LEVELS=3 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py -x -q -m gpu > gpurun_out/r2u_pytest_tl3.log 2>&1; echo "pytest(TL=3) rc=$?"; tail -3 gpurun_out/r2u_pytest_tl3.log
