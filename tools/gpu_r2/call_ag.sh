#!/bin/bash
# round 2, call AG: after SINGLE tiles -- new parity test, fuzz soak x8, ncu --set full of configs 2 and 3, launch list of the bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "tile_modes" > gpurun_out/r2ag_pytest_modes.log 2>&1; echo "pytest(tile modes) rc=$?"; tail -3 gpurun_out/r2ag_pytest_modes.log
TAHOE_FUZZ_SOAK=8 timeout 900 python -m pytest tests/test_gpu_fuzz.py -x -q -m gpu > gpurun_out/r2ag_fuzz.log 2>&1; echo "fuzz soak rc=$?"; tail -2 gpurun_out/r2ag_fuzz.log
cap() {  # cfg strategy kernel-regex
  python tools/run_one.py $1 $2 0 3 > gpurun_out/r2ag_plain_$1.log 2>&1 || { echo "$1 plain run failed"; tail -3 gpurun_out/r2ag_plain_$1.log; return; }
  tail -1 gpurun_out/r2ag_plain_$1.log
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$3 -s 1 -c 1 -f -o gpurun_out/r2ag_full_$1 \
      python tools/run_one.py $1 $2 0 3 > gpurun_out/r2ag_ncu_$1.log 2>&1
  echo "ncu $1 rc=$?"
}
cap C3 4 k_smem_forest
cap C2 4 k_smem_forest
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-configs > gpurun_out/r2ag_bench_plain.json 2> gpurun_out/r2ag_bench_plain.err || { echo "plain bench failed"; tail -5 gpurun_out/r2ag_bench_plain.err; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2ag_launches_bench_C3.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-configs > gpurun_out/r2ag_ncu_launches.log 2>&1
echo "launch list rc=$?"; grep -c k_smem_forest gpurun_out/r2ag_launches_bench_C3.csv
