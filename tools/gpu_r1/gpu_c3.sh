#!/bin/bash
mkdir -p gpurun_out; L=gpurun_out/c3_layout.log; : > $L
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q > gpurun_out/pytest_parity.log 2>&1; tail -3 gpurun_out/pytest_parity.log
timeout 300 python tools/probe_prof.py C3 2000000 >> $L 2>&1
timeout 300 python tools/probe_prof.py C2 >> $L 2>&1
TAHOE_SMEM_PLAN=16,1,6 timeout 300 python tools/probe_prof.py C2 >> $L 2>&1
TAHOE_SMEM_PLAN=8,1,6 timeout 300 python tools/probe_prof.py C3 2000000 >> $L 2>&1
TAHOE_SMEM_PLAN=16,1,4 timeout 300 python tools/probe_prof.py C3 2000000 >> $L 2>&1
timeout 600 python tools/probe_prof.py C3 10000000 >> $L 2>&1
cut -c47-300 $L
