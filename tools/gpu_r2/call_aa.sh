#!/bin/bash
# round 2, call AA: after the SHARED_DATA rewrite -- model check (config-1 sweep, hold-out sweep), both bench arms
mkdir -p gpurun_out
timeout 600 python tools/sweep.py C1 200000 > gpurun_out/r2aa_sweep_C1.json 2> gpurun_out/r2aa_sweep_C1.log; echo "sweep C1 rc=$?"; tail -12 gpurun_out/r2aa_sweep_C1.log
timeout 900 python tools/sweep_holdout.py > gpurun_out/r2aa_sweep_holdout.json 2> gpurun_out/r2aa_sweep_holdout.log; echo "holdout rc=$?"; grep -c "slowdown=1.00\|slow=1.00" gpurun_out/r2aa_sweep_holdout.log; tail -3 gpurun_out/r2aa_sweep_holdout.log
(time timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2aa_bench_ref.json 2> gpurun_out/r2aa_bench_ref.err); echo "ref rc=$?"
(time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2aa_bench.json 2> gpurun_out/r2aa_bench.err); echo "bench rc=$?"; head -c 300 gpurun_out/r2aa_bench.json; echo
