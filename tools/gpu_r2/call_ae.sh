#!/bin/bash
# round 2, call AE: SINGLE tiles (one TMA buffer per warp, next tile prefetched into L2) against staged / double-buffered plans
mkdir -p gpurun_out
{
for plan in auto 16,2,5 16,2,6 12,2,5; do TAHOE_SMEM_PLAN=$plan python tools/probe_prof.py C2 2>&1 | tail -1; done
for plan in auto 12,2,5 12,2,4 8,2,6 16,2,5 16,2,4; do TAHOE_SMEM_PLAN=$plan python tools/probe_prof.py C3 2000000 2>&1 | tail -1; done
} | tee gpurun_out/r2ae_single.log
