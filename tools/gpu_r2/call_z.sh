#!/bin/bash
# round 2, call Z2: SHARED_DATA final form (walk's own node word, next row prefetched into L2, tight top image by bulk copy,
# plan = fewest L2 trips with >= 10 teams): config-1 plans, full GPU suite, ncu --set full of the kernel
mkdir -p gpurun_out
{
python tools/probe_c1.py
TAHOE_SD_TOP_LEVELS=5 python tools/probe_c1.py
python tools/probe_c1.py 200000
python tools/probe_c1.py 10000
python tools/probe_c1.py 1000
TAHOE_SD_TOP_LEVELS=5 python tools/probe_c1.py 1000
} 2>&1 | grep "^C1" | tee gpurun_out/r2z_c1.log
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/r2z_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2z_pytest.log
python tools/run_one.py C1 2 0 3 2>&1 | tail -1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_shared_data -s 1 -c 1 -f -o gpurun_out/r2z_full_C1 \
   python tools/run_one.py C1 2 0 3 > gpurun_out/r2z_ncu_C1.log 2>&1; echo "ncu rc=$?"
