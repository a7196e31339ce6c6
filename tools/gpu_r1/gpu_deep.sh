#!/bin/bash
# k_smem_forest_deep (pipelined deep walks): sparse parity tests, then depth x K sweep on config 4
mkdir -p gpurun_out; L=gpurun_out/deep.log; : > $L
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "sparse" > gpurun_out/pytest_deep.log 2>&1; tail -5 gpurun_out/pytest_deep.log
for d in ${DEEP_DEPTHS:-5 6 7 8 9}; do
  TAHOE_QTOP_DEPTH=$d timeout 300 python tools/probe_qtop.py 1000000 >> $L 2>&1
done
for k in ${DEEP_KS:-2 6 8}; do
  echo "== K=$k" >> $L
  for d in ${DEEP_KDEPTHS:-6 7 8}; do
    TAHOE_B200_LIB=tahoe_b200/lib/libtahoe_b200_k$k.so TAHOE_QTOP_DEPTH=$d timeout 300 python tools/probe_qtop.py 1000000 >> $L 2>&1
  done
done
echo "== no drain" >> $L
for d in 6 8; do TAHOE_QTOP_NODRAIN=1 TAHOE_QTOP_DEPTH=$d timeout 300 python tools/probe_qtop.py 1000000 >> $L 2>&1; done
cut -c1-40,100-400 $L
