#!/bin/bash
# round 2, call AI: the final library -- full GPU suite, smoke, both bench arms
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/r2ai_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2ai_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
(time timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2ai_bench_ref.json 2> gpurun_out/r2ai_bench_ref.err); echo "ref rc=$?"
(time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2ai_bench.json 2> gpurun_out/r2ai_bench.err); echo "bench rc=$?"; head -c 250 gpurun_out/r2ai_bench.json; echo
