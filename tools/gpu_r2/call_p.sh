#!/bin/bash
# round 2, call P: quantise-from-tiles as the automatic choice for strided tables: suite (+ forced variant), soak, bench
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2p_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2p_pytest.log
TAHOE_QUANT_TILES=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py tests/test_gpu_full_size.py tests/test_zz_no_alloc.py -x -q > gpurun_out/r2p_pytest_tiles.log 2>&1; echo "pytest(TAHOE_QUANT_TILES=1) rc=$?"; tail -3 gpurun_out/r2p_pytest_tiles.log | cut -c1-200
(time timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2p_bench.json 2> gpurun_out/r2p_bench.err); echo "bench rc=$?"; head -c 250 gpurun_out/r2p_bench.json; echo
