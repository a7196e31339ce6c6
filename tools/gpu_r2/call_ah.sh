#!/bin/bash
# round 2, call AH: final library -- model sweeps (hold-out, C2, C3, C4), both bench arms
mkdir -p gpurun_out
timeout 900 python tools/sweep_holdout.py > gpurun_out/r2ah_sweep_holdout.json 2> gpurun_out/r2ah_sweep_holdout.log; echo "holdout rc=$?"; grep -v "slowdown=1.00" gpurun_out/r2ah_sweep_holdout.log | cut -c1-220
for c in "C2 100000000" "C3 100000000" "C4 1000000" "C1 10000000"; do set -- $c
  timeout 900 python tools/sweep.py $1 $2 > gpurun_out/r2ah_sweep_$1.json 2> gpurun_out/r2ah_sweep_$1.log; echo "sweep $1 rc=$?"; grep -v "slowdown=1.00" gpurun_out/r2ah_sweep_$1.log | cut -c1-220
done
(time timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2ah_bench_ref.json 2> gpurun_out/r2ah_bench_ref.err); echo "ref rc=$?"
(time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2ah_bench.json 2> gpurun_out/r2ah_bench.err); echo "bench rc=$?"; head -c 300 gpurun_out/r2ah_bench.json; echo
