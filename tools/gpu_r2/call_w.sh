#!/bin/bash
# round 2, call W: SHARED_DATA with a 64-byte tail block (two decisions + the leaf in one L2 trip): config 1 plans, parity
mkdir -p gpurun_out
{
TAHOE_SD_NO_TAIL=1 python tools/probe_c1.py
python tools/probe_c1.py
TAHOE_SD_TOP_LEVELS=4 python tools/probe_c1.py
TAHOE_SD_TOP_LEVELS=5 python tools/probe_c1.py
TAHOE_SD_TOP_LEVELS=6 python tools/probe_c1.py
TAHOE_SD_TOP_LEVELS=6 TAHOE_SD_TEAM=1 python tools/probe_c1.py
python tools/probe_c1.py 200000
} 2>&1 | grep "^C1" | tee gpurun_out/r2w_c1.log
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py -x -q -m gpu > gpurun_out/r2w_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2w_pytest.log
