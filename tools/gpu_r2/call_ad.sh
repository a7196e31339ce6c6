#!/bin/bash
# round 2, call AD: k_smem_forest_pair on very deep dense trees (few trees per chunk: few walks per thread)
mkdir -p gpurun_out
{
for pr in 0 1; do
  TAHOE_PAIR=$pr python tools/probe_deep_dense.py 200 13 28 1000000 2>&1 | tail -1
  TAHOE_PAIR=$pr TAHOE_SMEM_PLAN=16,0,2 python tools/probe_deep_dense.py 200 13 28 1000000 2>&1 | tail -1
  TAHOE_PAIR=$pr TAHOE_SMEM_PLAN=8,0,3 python tools/probe_deep_dense.py 200 13 28 1000000 2>&1 | tail -1
  TAHOE_PAIR=$pr python tools/probe_deep_dense.py 100 14 28 1000000 2>&1 | tail -1
  TAHOE_PAIR=$pr TAHOE_SMEM_PLAN=16,0,1 python tools/probe_deep_dense.py 100 14 28 1000000 2>&1 | tail -1
done
} | tee gpurun_out/r2ad_pair_deep.log
