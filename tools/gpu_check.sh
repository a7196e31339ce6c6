#!/bin/bash
# state check after a restore: full GPU suite, bench line, QT probe on config 4
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_chk.json 2> gpurun_out/bench_chk.err; echo "bench rc=$?"; cat gpurun_out/bench_chk.json
timeout 300 python tools/probe_qtop.py 1000000 2>&1 | tail -2
timeout 300 python tools/probe_sparse.py 1000000 2>&1 | tail -1
