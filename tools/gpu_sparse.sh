#!/bin/bash
mkdir -p gpurun_out; L=gpurun_out/sparse_smem.log; : > $L
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "sparse" > gpurun_out/pytest_sparse.log 2>&1; tail -15 gpurun_out/pytest_sparse.log
timeout 300 python tools/probe_sparse.py 200000 >> $L 2>&1
for cut in 64 128 256 512 1024; do
  TAHOE_SPARSE_CUT=$cut timeout 300 python tools/probe_sparse.py 1000000 >> $L 2>&1
done
timeout 200 python tools/run_one.py C4 0 1000000 2 >> $L 2>&1
cat $L
