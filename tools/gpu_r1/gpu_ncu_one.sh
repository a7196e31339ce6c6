#!/bin/bash
# tools/gpu_ncu_one.sh TAG KERNEL_REGEX SKIP -- ARGS...   one full ncu capture of run_one.py ARGS (plain run first)
TAG=$1; KRE=$2; SKIP=$3; shift 4
mkdir -p gpurun_out
python tools/run_one.py "$@" > gpurun_out/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_$TAG.log; exit 1; }
cat gpurun_out/plain_$TAG.log
ncu --set full --clock-control none --import-source on -k regex:"$KRE" -s $SKIP -c 1 -f -o gpurun_out/prof_$TAG \
    python tools/run_one.py "$@" > gpurun_out/ncu_$TAG.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_$TAG.log
