#!/bin/bash
# tools/gpu_plans.sh -- per-kernel timings of shared-memory-forest plans
# (TAHOE_SMEM_PLAN="warps,staged,group") on C2 (5M rows) and C3 (2M rows).
mkdir -p gpurun_out
: > gpurun_out/plans.log
for plan in auto 16,0,6 16,1,6 16,1,5 16,1,4 16,1,3 8,1,6 8,0,6 16,0,4; do
  if [ $plan = auto ]; then unset TAHOE_SMEM_PLAN; else export TAHOE_SMEM_PLAN=$plan; fi
  timeout 200 python tools/probe_prof.py C2 >> gpurun_out/plans.log 2>&1
done
for plan in auto 16,1,5 16,1,4 16,1,3 16,1,2 8,1,6 8,1,5 8,1,4 8,1,3 8,1,2 4,1,6 4,1,3 16,0,3 8,0,5; do
  if [ $plan = auto ]; then unset TAHOE_SMEM_PLAN; else export TAHOE_SMEM_PLAN=$plan; fi
  timeout 200 python tools/probe_prof.py C3 2000000 >> gpurun_out/plans.log 2>&1
done
unset TAHOE_SMEM_PLAN
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k sparse 2>&1 | tail -2
cut -c47-300 gpurun_out/plans.log
