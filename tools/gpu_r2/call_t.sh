#!/bin/bash
# round 2, call T: final library (plain kernel default, quantise from tiles where the table is strided): suite, smoke, both bench arms, launch list
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/r2t_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2t_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
(time timeout 600 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2t_bench_ref.json 2> gpurun_out/r2t_bench_ref.err); echo "ref rc=$?"
(time timeout 900 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2t_bench.json 2> gpurun_out/r2t_bench.err); echo "bench rc=$?"; head -c 250 gpurun_out/r2t_bench.json; echo
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2t_launches_bench_C3.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-configs > gpurun_out/r2t_ncu_launches.log 2>&1; echo "launch list rc=$?"
