#!/bin/bash
# round 2, call E: recalibrated plans (C2 16,1,5; C3 12,0,4; C1 2-warp single-buffer teams): suite, probes, bench, hold-out sweep
mkdir -p gpurun_out; L=gpurun_out/r2e_plans.log; : > $L
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2e_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2e_pytest.log
timeout 300 python tools/probe_prof.py C2 >> $L 2>&1
timeout 300 python tools/probe_prof.py C3 2000000 >> $L 2>&1
timeout 300 python tools/probe_prof.py C3 1250000 >> $L 2>&1
timeout 300 python tools/probe_prof.py C3 10000000 >> $L 2>&1
TAHOE_SMEM_PLAN=8,1,6 timeout 300 python tools/probe_prof.py C3 10000000 >> $L 2>&1
TAHOE_SMEM_PLAN=8,1,6 timeout 300 python tools/probe_prof.py C3 1250000 >> $L 2>&1
TAHOE_SMEM_PLAN=12,0,4 TAHOE_TOP_LEVELS=2 timeout 300 python tools/probe_prof.py C3 10000000 >> $L 2>&1
cut -c1-215 $L | sed "s/checksum.*//"
timeout 200 python tools/probe_c1.py 2>&1 | tail -1
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err; echo "bench rc=$?"; tail -c 300 gpurun_out/r2e_bench.err; head -c 400 gpurun_out/r2e_bench.json; echo
timeout 1500 python tools/sweep_holdout.py 10000000 > gpurun_out/r2e_sweep_holdout.json 2> gpurun_out/r2e_sweep_holdout.log; echo "holdout rc=$?"; grep "^#" gpurun_out/r2e_sweep_holdout.log | cut -c1-220
