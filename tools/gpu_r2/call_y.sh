#!/bin/bash
# round 2, call Y3: SHARED_DATA after the walk rewrite + L2 prefetch of the next row + early first load: plans, full GPU suite
mkdir -p gpurun_out
{
python tools/probe_c1.py
TAHOE_SD_NO_L2_AHEAD=1 python tools/probe_c1.py
TAHOE_SD_TOP_LEVELS=3 python tools/probe_c1.py
TAHOE_SD_TOP_LEVELS=7 python tools/probe_c1.py
TAHOE_SD_TEAM=1 python tools/probe_c1.py
python tools/probe_c1.py 200000
python tools/probe_c1.py 1000
PROBE_TREES=1 PROBE_DEPTH=1 TAHOE_SD_BUFS=1 python tools/probe_c1.py
} 2>&1 | grep "^C1" | tee gpurun_out/r2y_c1.log
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/r2y_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2y_pytest.log
