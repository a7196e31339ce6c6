#!/bin/bash
mkdir -p gpurun_out; L=gpurun_out/r2s_hostdev.log; : > $L
TAHOE_QUANT_TILES=1 TAHOE_CLUSTER=0 timeout 300 python tools/probe_host_vs_device.py C3 10000000 >> $L 2>&1
TAHOE_QUANT_TILES=1 TAHOE_CLUSTER=0 TAHOE_SMEM_PLAN=8,1,6 timeout 300 python tools/probe_host_vs_device.py C3 10000000 >> $L 2>&1
TAHOE_QUANT_TILES=1 timeout 300 python tools/probe_host_vs_device.py C2 5000000 >> $L 2>&1
TAHOE_QUANT_TILES=1 TAHOE_CLUSTER=2 timeout 300 python tools/probe_host_vs_device.py C3 10000000 >> $L 2>&1
cat $L | cut -c1-400
