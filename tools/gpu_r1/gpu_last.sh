#!/bin/bash
# last pass of the round: smoke, full GPU suite, bench (both arms)
mkdir -p gpurun_out
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_last.json 2> gpurun_out/bench_last.err; echo "bench rc=$?"; cat gpurun_out/bench_last.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_last.json 2> gpurun_out/bench_ref_last.err; echo "ref rc=$?"; cat gpurun_out/bench_ref_last.json
