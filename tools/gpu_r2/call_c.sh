#!/bin/bash
# round 2, call C: no-alloc entry point alone (fresh process), full suite, forced split launches, plan probes
mkdir -p gpurun_out; L=gpurun_out/r2c_plans.log; : > $L
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/r2c_pytest_noalloc.log 2>&1; echo "pytest(no_alloc alone) rc=$?"; tail -4 gpurun_out/r2c_pytest_noalloc.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/r2c_pytest.log
TAHOE_TOP_TREES=12 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py -x -q > gpurun_out/r2c_pytest_split.log 2>&1; echo "pytest(split launches) rc=$?"; tail -3 gpurun_out/r2c_pytest_split.log
for tl in 0 1 2; do
  TAHOE_TOP_LEVELS=$tl TAHOE_SMEM_PLAN=8,1,6 timeout 300 python tools/probe_prof.py C3 2000000 >> $L 2>&1
  TAHOE_TOP_LEVELS=$tl TAHOE_SMEM_PLAN=12,1,5 timeout 300 python tools/probe_prof.py C3 2000000 >> $L 2>&1
  TAHOE_TOP_LEVELS=$tl TAHOE_SMEM_PLAN=16,1,5 timeout 300 python tools/probe_prof.py C3 2000000 >> $L 2>&1
  TAHOE_TOP_LEVELS=$tl timeout 300 python tools/probe_prof.py C2 >> $L 2>&1
  TAHOE_TOP_LEVELS=$tl TAHOE_SMEM_PLAN=16,1,5 timeout 300 python tools/probe_prof.py C2 >> $L 2>&1
done
TAHOE_SMEM_PLAN=12,0,4 timeout 300 python tools/probe_prof.py C3 2000000 >> $L 2>&1
TAHOE_SMEM_PLAN=12,1,5 timeout 300 python tools/probe_prof.py C2 >> $L 2>&1
timeout 300 python tools/probe_qtop.py 1000000 >> $L 2>&1
cut -c1-230 $L | sed "s/checksum.*//"
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
