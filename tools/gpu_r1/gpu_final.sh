#!/bin/bash
# end-of-round pass: full GPU suite, bench (+ reference arm), ncu artefacts of the bench command, C3/C4 checks
TAG=${1:-r01c}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo "bench rc=$?"; cat gpurun_out/bench_$TAG.json
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2> gpurun_out/bench_ref_$TAG.err; echo "ref rc=$?"; cat gpurun_out/bench_ref_$TAG.json
bash tools/gpu_r1/gpu_ncu.sh $TAG
timeout 600 python tools/probe_prof.py C3 10000000 2>&1 | cut -c47-300
timeout 300 python tools/run_one.py C1 2 0 4 2>&1 | tail -2
timeout 300 python tools/probe_sparse.py 1000000 2>&1 | tail -1
timeout 900 python tools/sweep.py C3 10000000 > gpurun_out/sweep_C3.json 2> gpurun_out/sweep_C3.log; echo "C3 sweep rc=$?"
timeout 600 python tools/sweep.py C4 1000000 > gpurun_out/sweep_C4.json 2> gpurun_out/sweep_C4.log; echo "C4 sweep rc=$?"
grep "^#" gpurun_out/sweep_C3.log | cut -c1-120; grep "^#" gpurun_out/sweep_C4.log | cut -c1-120
