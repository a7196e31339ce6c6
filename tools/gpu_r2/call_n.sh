#!/bin/bash
# round 2, call N: fuzz soak x8 and the full suite on the final library
mkdir -p gpurun_out
TAHOE_FUZZ_SOAK=8 timeout 1500 python -m pytest tests/test_gpu_fuzz.py -q > gpurun_out/r2n_fuzz_soak.log 2>&1; echo "fuzz soak rc=$?"; tail -3 gpurun_out/r2n_fuzz_soak.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2n_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2n_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
