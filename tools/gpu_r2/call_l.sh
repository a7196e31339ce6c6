#!/bin/bash
# round 2, call L: fuzz soak on the final library (dense depths to 13 now reach the cluster pipeline), then the bench line
mkdir -p gpurun_out
TAHOE_FUZZ_SOAK=8 timeout 1500 python -m pytest tests/test_gpu_fuzz.py -x -q > gpurun_out/r2l_fuzz_soak.log 2>&1; echo "fuzz soak rc=$?"; tail -4 gpurun_out/r2l_fuzz_soak.log
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2l_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2l_pytest.log
(time timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2l_bench.json 2> gpurun_out/r2l_bench.err); echo "bench rc=$?"; head -c 250 gpurun_out/r2l_bench.json; echo
