#!/bin/bash
# tools/gpu_first.sh -- first GPU pass: parity tests, smoke, per-strategy probe, bench, launch list.
# Run under gpurun from the repo root; everything lands in gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
nproc > gpurun_out/nproc.txt; lscpu | head -20 >> gpurun_out/nproc.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/smoke.log
timeout 600 python tools/probe.py C2 > gpurun_out/probe_C2.log 2>&1; echo "probe rc=$?"
timeout 300 python tools/probe_prof.py C2 > gpurun_out/probe_prof_C2.log 2>&1
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "bench_ref rc=$?"
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/smoke.log | tail -3; cat gpurun_out/probe_C2.log; cat gpurun_out/probe_prof_C2.log; cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err; cat gpurun_out/bench_ref.json
