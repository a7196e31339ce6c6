#!/bin/bash
# sparse parity, default config-4 plan, one full ncu capture of k_smem_forest_deep, reference-GPU comparison
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "sparse" > gpurun_out/pytest_deep.log 2>&1; tail -3 gpurun_out/pytest_deep.log
timeout 300 python tools/probe_qtop.py 1000000 2>&1 | tail -1
timeout 300 python tools/run_qtop.py 1000000 3 > gpurun_out/plain_c4q.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_smem_forest_deep -s 1 -c 1 -f -o gpurun_out/prof_c4q \
    python tools/run_qtop.py 1000000 3 > gpurun_out/ncu_c4q.log 2>&1
echo "ncu rc=$?"; cat gpurun_out/plain_c4q.log; tail -2 gpurun_out/ncu_c4q.log
bash tools/gpu_ref_compare.sh
