#!/bin/bash
# QT path (sparse forests, quantised tops): parity tests, then depth / plan sweep on config 4
mkdir -p gpurun_out; L=gpurun_out/qtop.log; : > $L
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "sparse" > gpurun_out/pytest_qtop.log 2>&1; tail -5 gpurun_out/pytest_qtop.log
echo "== default plan" >> $L
timeout 300 python tools/probe_qtop.py 1000000 >> $L 2>&1
for d in ${QTOP_DEPTHS:-7 9 10 11}; do
  TAHOE_QTOP_DEPTH=$d timeout 300 python tools/probe_qtop.py 1000000 >> $L 2>&1
done
echo "== no drain (top part alone; scores wrong)" >> $L
for d in 6 8 10; do
  TAHOE_QTOP_NODRAIN=1 TAHOE_QTOP_DEPTH=$d timeout 300 python tools/probe_qtop.py 1000000 >> $L 2>&1
done
echo "== other plans" >> $L
for plan in ${QTOP_PLANS:-"8,0,6" "8,1,4"}; do
  for d in 8 9; do
    TAHOE_QTOP_DEPTH=$d TAHOE_SMEM_PLAN=$plan timeout 300 python tools/probe_qtop.py 1000000 >> $L 2>&1
  done
done
cat $L
