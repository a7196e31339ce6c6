#!/bin/bash
# call S (third form): ten repeated host-path calls per variant -- is the plain kernel clean, does the cluster kernel fail without from-tiles
mkdir -p gpurun_out; L=gpurun_out/r2s_hostdev10.log; : > $L
TAHOE_QUANT_TILES=1 TAHOE_CLUSTER=0 timeout 400 python tools/probe_host_vs_device.py C3 10000000 10 >> $L 2>&1
TAHOE_QUANT_TILES=0 TAHOE_CLUSTER=2 timeout 400 python tools/probe_host_vs_device.py C3 10000000 10 >> $L 2>&1
TAHOE_QUANT_TILES=1 TAHOE_CLUSTER=2 TAHOE_HOST_SLAB=30,100,30 timeout 400 python tools/probe_host_vs_device.py C3 10000000 6 >> $L 2>&1
timeout 400 python tools/probe_host_vs_device.py C2 5000000 10 >> $L 2>&1
grep -c "rows 0$" $L; grep -v "rows 0$" $L | cut -c1-330
