#!/bin/bash
# round 2, call AC: k_smem_forest_pair (two tiles per warp walked together) on config 3 against the plain kernel
mkdir -p gpurun_out
{
for pr in 0 1; do
  for plan in 12,0,4 8,0,5 16,0,3 8,0,4; do
    TAHOE_PAIR=$pr TAHOE_SMEM_PLAN=$plan python tools/probe_prof.py C3 2000000 2>&1 | tail -1 | sed "s/^/PAIR=$pr /"
  done
done
TAHOE_PAIR=1 python tools/probe_prof.py C3 2>&1 | tail -1 | sed "s/^/PAIR=1 /"
TAHOE_PAIR=0 python tools/probe_prof.py C3 2>&1 | tail -1 | sed "s/^/PAIR=0 /"
} | tee gpurun_out/r2ac_pair.log
