#!/bin/bash
# round 2, call B: top levels as kernel parameters -- suite, plans on C2/C3/C4, ncu of C3 (staged and TMA)
mkdir -p gpurun_out; L=gpurun_out/r2b_plans.log; : > $L
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2b_pytest.log
TAHOE_TOP_TREES=12 timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py -x -q > gpurun_out/r2b_pytest_split.log 2>&1; echo "pytest(split launches) rc=$?"; tail -3 gpurun_out/r2b_pytest_split.log
for plan in 8,1,6 8,0,5 16,0,3 16,1,5; do
  TAHOE_SMEM_PLAN=$plan timeout 300 python tools/probe_prof.py C3 2000000 >> $L 2>&1
done
timeout 300 python tools/probe_prof.py C2 >> $L 2>&1
TAHOE_SMEM_PLAN=16,1,5 timeout 300 python tools/probe_prof.py C2 >> $L 2>&1
timeout 300 python tools/probe_prof.py C3 10000000 >> $L 2>&1
timeout 300 python tools/probe_qtop.py 1000000 >> $L 2>&1
cut -c47-300 $L
for plan in 8,1,6 8,0,5; do
  tag=$(echo $plan | tr , _)
  TAHOE_SMEM_PLAN=$plan timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_smem_forest -s 1 -c 1 -f \
      -o gpurun_out/r2b_ncu_c3_$tag python tools/run_one.py C3 4 2000000 3 > gpurun_out/r2b_ncu_$tag.log 2>&1
  echo "ncu $plan rc=$?"
done
ls -la gpurun_out/*.ncu-rep
