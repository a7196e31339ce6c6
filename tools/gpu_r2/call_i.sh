#!/bin/bash
# round 2, call I: cluster pipeline as the default for config 3 -- suite, bench, full ncu capture of the cluster kernel
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2i_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2i_pytest.log
(time timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench.err); echo "bench rc=$?"; head -c 300 gpurun_out/r2i_bench.json; echo
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_smem_forest -s 1 -c 1 -f -o gpurun_out/r2i_full_C3_cluster \
      python tools/run_one.py C3 4 0 3 > gpurun_out/r2i_ncu_C3.log 2>&1; echo "ncu rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2i_launches_bench_C3.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-configs > gpurun_out/r2i_ncu_launches.log 2>&1; echo "launch list rc=$?"
