#!/bin/bash
# tools/gpu_variants.sh -- parity tests on the default build, then per-kernel timings of every
# variant library present (tools/build_variant.sh) on C2.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/pytest_gpu.log
: > gpurun_out/variants.log
timeout 200 python tools/probe_prof.py C2 >> gpurun_out/variants.log 2>&1
for lib in tahoe_b200/lib/libtahoe_b200_*.so; do
  TAHOE_B200_LIB=$PWD/$lib timeout 200 python tools/probe_prof.py C2 >> gpurun_out/variants.log 2>&1
done
cat gpurun_out/variants.log
