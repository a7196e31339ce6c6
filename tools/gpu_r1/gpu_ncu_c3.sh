#!/bin/bash
mkdir -p gpurun_out
python tools/run_one.py C3 4 2000000 3 > gpurun_out/plain_c3.log 2>&1 || { tail -5 gpurun_out/plain_c3.log; exit 1; }
cat gpurun_out/plain_c3.log
ncu --set full --clock-control none --import-source on -k regex:k_smem_forest -s 1 -c 1 -f -o gpurun_out/prof_c3 \
    python tools/run_one.py C3 4 2000000 3 > gpurun_out/ncu_c3.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_c3.log
