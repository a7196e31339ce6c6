#!/bin/bash
# round 2, call O: k_quantize from column-major float tiles (transpose first, quantise in place) vs from the rows
mkdir -p gpurun_out
TAHOE_QUANT_TILES=1 timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fuzz.py tests/test_gpu_full_size.py -x -q > gpurun_out/r2o_pytest_tiles.log 2>&1; echo "pytest(TAHOE_QUANT_TILES=1) rc=$?"; tail -4 gpurun_out/r2o_pytest_tiles.log | cut -c1-200
L=gpurun_out/r2o_quant.log; : > $L
for q in 0 1; do
  TAHOE_QUANT_TILES=$q timeout 200 python tools/probe_prof.py C2 >> $L 2>&1
  TAHOE_QUANT_TILES=$q timeout 200 python tools/probe_prof.py C3 10000000 >> $L 2>&1
  TAHOE_QUANT_TILES=$q timeout 200 python tools/probe_prof.py C3 1250000 >> $L 2>&1
  TAHOE_QUANT_TILES=$q timeout 300 python tools/probe_qtop.py 1000000 >> $L 2>&1
done
sed "s/checksum.*//" $L | cut -c1-260
