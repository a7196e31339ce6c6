#!/bin/bash
mkdir -p gpurun_out; L=gpurun_out/sparse_smem.log; : > $L
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "sparse" > gpurun_out/pytest_sparse.log 2>&1; tail -3 gpurun_out/pytest_sparse.log
for g in 2 3 4 1; do
  echo "== group $g" >> $L
  TAHOE_SPARSE_GROUP=$g timeout 300 python tools/probe_sparse.py 1000000 >> $L 2>&1
done
cat $L
