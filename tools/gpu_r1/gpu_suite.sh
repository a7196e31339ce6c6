#!/bin/bash
# full GPU suite + config-4 batch sweep (performance-model check)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 900 python tools/sweep.py C4 1000000 > gpurun_out/sweep_C4.json 2> gpurun_out/sweep_C4.log; echo "C4 sweep rc=$?"
grep "^#" gpurun_out/sweep_C4.log | cut -c1-160
