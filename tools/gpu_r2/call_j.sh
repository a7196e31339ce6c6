#!/bin/bash
# round 2, call J: first tile of a pass fetched before the pass barrier -- suite, then the probes that show fixed per-pass cost
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2j_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2j_pytest.log
L=gpurun_out/r2j_plans.log; : > $L
timeout 200 python tools/probe_prof.py C3 1250000 >> $L 2>&1
timeout 200 python tools/probe_prof.py C3 10000000 >> $L 2>&1
TAHOE_CLUSTER=0 timeout 200 python tools/probe_prof.py C3 1250000 >> $L 2>&1
timeout 200 python tools/probe_prof.py C2 >> $L 2>&1
timeout 200 python tools/probe_prof.py C2 625000 >> $L 2>&1
timeout 300 python tools/probe_qtop.py 1000000 >> $L 2>&1
sed "s/checksum.*//" $L | cut -c1-215
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r2j_bench.json 2> gpurun_out/r2j_bench.err; echo "bench rc=$?"; head -c 200 gpurun_out/r2j_bench.json; echo
