#!/bin/bash
# round 2: ncu evidence for the final kernels.  (1) launch list of the bench command (per-launch durations:
# the kernels' SHARE of the step); (2) one --set full capture of the dominant kernel of every config at its
# full size.  Everything lands in gpurun_out/; tools/ncu_summary.py turns the reports into profiles/r02_*.json.
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-configs > gpurun_out/r2n_bench_plain.json 2> gpurun_out/r2n_bench_plain.err || { echo "plain bench failed"; tail -5 gpurun_out/r2n_bench_plain.err; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2n_launches_bench_C3.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --no-configs > gpurun_out/r2n_ncu_launches.log 2>&1
echo "launch list rc=$?"; grep -c k_smem_forest gpurun_out/r2n_launches_bench_C3.csv
cap() {  # cfg strategy kernel-regex
  python tools/run_one.py $1 $2 0 3 > gpurun_out/r2n_plain_$1.log 2>&1 || { echo "$1 plain run failed"; tail -3 gpurun_out/r2n_plain_$1.log; return; }
  tail -1 gpurun_out/r2n_plain_$1.log
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$3 -s 1 -c 1 -f -o gpurun_out/r2n_full_$1 \
      python tools/run_one.py $1 $2 0 3 > gpurun_out/r2n_ncu_$1.log 2>&1
  echo "ncu $1 rc=$?"
}
cap C3 4 k_smem_forest
cap C2 4 k_smem_forest
cap C4 4 k_smem_forest_deep
cap C1 2 k_shared_data
timeout 600 ncu --set full --clock-control none -k regex:k_quantize -s 1 -c 1 -f -o gpurun_out/r2n_full_C4_quantize \
    python tools/run_one.py C4 4 0 3 > gpurun_out/r2n_ncu_C4q.log 2>&1; echo "ncu C4 quantize rc=$?"
ls -la gpurun_out/*.ncu-rep
