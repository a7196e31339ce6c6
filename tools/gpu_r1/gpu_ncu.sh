#!/bin/bash
# tools/gpu_ncu.sh [TAG] -- launch list + one full ncu capture of the hot kernels for the bench command.
# The plain run goes first (same command line), ncu only if it exits 0.
TAG=${1:-r01}
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv \
    --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launches_$TAG.log 2>&1
echo "launch list rc=$?"
$CMD > gpurun_out/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_smem_forest|k_quantize' -s 6 -c 2 \
    -f -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "full rc=$?"
cat gpurun_out/plain_$TAG.log | tail -2
tail -5 gpurun_out/ncu_full_$TAG.log
