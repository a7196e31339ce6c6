#!/bin/bash
# round 2, call X: SHARED_DATA with the walk's own node word (14 instructions per decision instead of ~25)
mkdir -p gpurun_out
{
python tools/probe_c1.py
TAHOE_SD_TOP_LEVELS=3 python tools/probe_c1.py
TAHOE_SD_TOP_LEVELS=4 python tools/probe_c1.py
TAHOE_SD_TOP_LEVELS=5 python tools/probe_c1.py
TAHOE_SD_TOP_LEVELS=6 python tools/probe_c1.py
TAHOE_SD_TOP_LEVELS=7 python tools/probe_c1.py
TAHOE_SD_TOP_LEVELS=5 TAHOE_SD_TEAM=1 python tools/probe_c1.py
TAHOE_SD_TOP_LEVELS=5 TAHOE_SD_BUFS=2 python tools/probe_c1.py
python tools/probe_c1.py 200000
} 2>&1 | grep "^C1" | tee gpurun_out/r2x_c1.log
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py tests/test_importers.py -x -q -m gpu > gpurun_out/r2x_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2x_pytest.log
