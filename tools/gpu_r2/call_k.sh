#!/bin/bash
# round 2, call K: back on the call-I kernels -- probes; the long batch sweeps (config 5): C3 to 10^8 rows, C1 to 10^7, C2 to 10^8
mkdir -p gpurun_out
L=gpurun_out/r2k_plans.log; : > $L
timeout 200 python tools/probe_prof.py C3 10000000 >> $L 2>&1
timeout 200 python tools/probe_prof.py C3 1250000 >> $L 2>&1
sed "s/checksum.*//" $L | cut -c1-215
timeout 1500 python tools/sweep.py C3 100000000 > gpurun_out/r2k_sweep_C3.json 2> gpurun_out/r2k_sweep_C3.log; echo "C3 sweep rc=$?"; grep "^#" gpurun_out/r2k_sweep_C3.log | cut -c1-200
timeout 900 python tools/sweep.py C1 10000000 > gpurun_out/r2k_sweep_C1.json 2> gpurun_out/r2k_sweep_C1.log; echo "C1 sweep rc=$?"; grep "^#" gpurun_out/r2k_sweep_C1.log | cut -c1-200
timeout 900 python tools/sweep.py C2 100000000 > gpurun_out/r2k_sweep_C2.json 2> gpurun_out/r2k_sweep_C2.log; echo "C2 sweep rc=$?"; grep "^#" gpurun_out/r2k_sweep_C2.log | cut -c1-200
timeout 600 python tools/sweep.py C4 1000000 > gpurun_out/r2k_sweep_C4.json 2> gpurun_out/r2k_sweep_C4.log; echo "C4 sweep rc=$?"; grep "^#" gpurun_out/r2k_sweep_C4.log | cut -c1-200
