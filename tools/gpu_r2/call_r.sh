#!/bin/bash
# round 2, call R: bench line with the per-config end-to-end legs
mkdir -p gpurun_out
(time timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2r_bench.json 2> gpurun_out/r2r_bench.err); echo "bench rc=$?"; tail -c 300 gpurun_out/r2r_bench.err; head -c 200 gpurun_out/r2r_bench.json; echo
