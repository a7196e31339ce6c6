#!/bin/bash
# round 2, call AF: SINGLE tiles chosen by the planner: auto plans at full size, full GPU suite, fuzz soak
mkdir -p gpurun_out
{
python tools/probe_prof.py C2 2>&1 | tail -1
python tools/probe_prof.py C3 2>&1 | tail -1
TAHOE_SMEM_PLAN=16,2,5 TAHOE_TOP_LEVELS=0 python tools/probe_prof.py C3 2>&1 | tail -1
TAHOE_SMEM_PLAN=12,0,4 python tools/probe_prof.py C3 2>&1 | tail -1
} | tee gpurun_out/r2af_single_full.log
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/r2af_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2af_pytest.log
for s in 1 2 3; do TAHOE_FUZZ_SEED_BASE=$((s*1000)) timeout 600 python -m pytest tests/test_gpu_fuzz.py -x -q -m gpu 2>&1 | tail -1; done
