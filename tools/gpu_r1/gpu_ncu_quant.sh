#!/bin/bash
# one full ncu capture of k_quantize on config 4 (strided-table path)
mkdir -p gpurun_out
timeout 300 python tools/run_qtop.py 1000000 3 > gpurun_out/plain_c4quant.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_quantize -s 1 -c 1 -f -o gpurun_out/prof_c4quant \
    python tools/run_qtop.py 1000000 3 > gpurun_out/ncu_c4quant.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_c4quant.log
