#!/bin/bash
# config 4: full GPU suite, probe, one full ncu capture of k_smem_forest_deep, batch sweep
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python tools/probe_qtop.py 1000000 2>&1 | tail -1
timeout 300 python tools/run_qtop.py 1000000 3 > gpurun_out/plain_c4q.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_smem_forest_deep -s 1 -c 1 -f -o gpurun_out/prof_c4q \
    python tools/run_qtop.py 1000000 3 > gpurun_out/ncu_c4q.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/plain_c4q.log; tail -2 gpurun_out/ncu_c4q.log
timeout 900 python tools/sweep.py C4 1000000 > gpurun_out/sweep_C4.json 2> gpurun_out/sweep_C4.log; echo "C4 sweep rc=$?"
grep "^#" gpurun_out/sweep_C4.log | cut -c1-170
