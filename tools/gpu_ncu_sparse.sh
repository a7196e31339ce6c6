#!/bin/bash
mkdir -p gpurun_out
python tools/run_sparse.py 300000 3 > gpurun_out/plain_c4s.log 2>&1 || { tail -5 gpurun_out/plain_c4s.log; exit 1; }
cat gpurun_out/plain_c4s.log
ncu --set full --clock-control none --import-source on -k regex:k_sparse_smem -s 1 -c 1 -f -o gpurun_out/prof_c4s \
    python tools/run_sparse.py 300000 3 > gpurun_out/ncu_c4s.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_c4s.log
