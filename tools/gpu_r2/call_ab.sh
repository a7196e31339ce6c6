#!/bin/bash
# round 2, call AB: the model after "latency + throughput" for the thread-per-row strategies: all sweeps again
mkdir -p gpurun_out
timeout 900 python tools/sweep_holdout.py > gpurun_out/r2ab_sweep_holdout.json 2> gpurun_out/r2ab_sweep_holdout.log; echo "holdout rc=$?"; grep -v "slowdown=1.00" gpurun_out/r2ab_sweep_holdout.log | cut -c1-200
for c in "C1 10000000" "C2 100000000" "C3 100000000" "C4 1000000"; do set -- $c
  timeout 900 python tools/sweep.py $1 $2 > gpurun_out/r2ab_sweep_$1.json 2> gpurun_out/r2ab_sweep_$1.log; echo "sweep $1 rc=$?"; grep -v "slowdown=1.00" gpurun_out/r2ab_sweep_$1.log | cut -c1-200
done
