#!/bin/bash
# round 2, call V: config 1's streaming ceiling -- the same rows under lighter and lighter walks
mkdir -p gpurun_out
{
python tools/probe_c1.py
PROBE_DEPTH=4 python tools/probe_c1.py
PROBE_DEPTH=2 python tools/probe_c1.py
PROBE_TREES=64 python tools/probe_c1.py
PROBE_TREES=64 PROBE_DEPTH=4 python tools/probe_c1.py
PROBE_TREES=32 PROBE_DEPTH=8 TAHOE_SD_BUFS=1 python tools/probe_c1.py
PROBE_TREES=1 PROBE_DEPTH=1 TAHOE_SD_BUFS=1 python tools/probe_c1.py
PROBE_TREES=1 PROBE_DEPTH=1 python tools/probe_c1.py
python tools/probe_c1.py 200000
PROBE_TREES=1 PROBE_DEPTH=1 TAHOE_SD_BUFS=1 python tools/probe_c1.py 200000
} 2>&1 | grep "^C1" | tee gpurun_out/r2v_c1.log
python -c "
from tahoe_b200 import api
print(api.measure_peaks())"
