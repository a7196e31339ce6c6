#!/bin/bash
# batch-size sweeps (BASELINE config 5) on all four forests + the full GPU test suite
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python tools/sweep.py C1 100000 > gpurun_out/sweep_C1.json 2> gpurun_out/sweep_C1.log; echo "C1 rc=$?"
timeout 600 python tools/sweep.py C2 100000000 > gpurun_out/sweep_C2.json 2> gpurun_out/sweep_C2.log; echo "C2 rc=$?"
timeout 900 python tools/sweep.py C3 10000000 > gpurun_out/sweep_C3.json 2> gpurun_out/sweep_C3.log; echo "C3 rc=$?"
timeout 600 python tools/sweep.py C4 1000000 > gpurun_out/sweep_C4.json 2> gpurun_out/sweep_C4.log; echo "C4 rc=$?"
for c in C1 C2 C3 C4; do python - <<PY
import json
try:
    d=json.load(open("gpurun_out/sweep_$c.json")); print("$c hit", round(d["hit_rate"],3), "worst", round(d["worst_pick_slowdown"],3))
except Exception as e: print("$c", e)
PY
done
