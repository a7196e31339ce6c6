#!/bin/bash
# round 2, call G: the cluster pipeline -- parity first (short timeouts: a protocol slip is a hang), then timing
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k cluster > gpurun_out/r2g_pytest_cluster.log 2>&1; rc=$?; echo "pytest(cluster) rc=$rc"; tail -15 gpurun_out/r2g_pytest_cluster.log | cut -c1-200
if [ $rc -ne 0 ]; then exit 0; fi
L=gpurun_out/r2g_cluster.log; : > $L
for S in 0 2 4 8; do
  TAHOE_CLUSTER=$S timeout 120 python tools/probe_prof.py C3 2000000 >> $L 2>&1
done
for S in 0 2 4; do
  TAHOE_CLUSTER=$S timeout 200 python tools/probe_prof.py C3 10000000 >> $L 2>&1
  TAHOE_CLUSTER=$S timeout 120 python tools/probe_prof.py C3 1250000 >> $L 2>&1
done
sed "s/checksum.*//" $L | awk -v OFS=" " '{print}' | cut -c1-220
grep -c . $L
