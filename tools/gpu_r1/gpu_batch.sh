#!/bin/bash
mkdir -p gpurun_out; : > gpurun_out/batch.log
for b in 0 16000000 32000000 48000000 64000000 96000000; do
  export TAHOE_ROW_BATCH_BYTES=$b
  echo "batch_bytes=$b" >> gpurun_out/batch.log
  timeout 200 python tools/probe_prof.py C2 >> gpurun_out/batch.log 2>&1
  timeout 200 python tools/probe_prof.py C3 2000000 >> gpurun_out/batch.log 2>&1
done
unset TAHOE_ROW_BATCH_BYTES
timeout 300 python tools/probe_prof.py C3 10000000 >> gpurun_out/batch.log 2>&1
cut -c1-40,60-250 gpurun_out/batch.log
