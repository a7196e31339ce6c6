#!/bin/bash
# round 2, call AJ: small-batch latency of SHARED_DATA: host launch path against kernel time (after caching the launch state)
mkdir -p gpurun_out
python tools/probe_small_batch.py 2>&1 | grep " R=\|Error\|error" | tee gpurun_out/r2aj_small2.log
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -2
