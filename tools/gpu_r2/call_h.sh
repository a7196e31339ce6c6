#!/bin/bash
# round 2, call H: full suite + smoke + bench on the final library; cluster S=2 vs none, interleaved on one box; DRAM bytes of S=2
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2h_pytest.log
L=gpurun_out/r2h_cluster_ab.log; : > $L
for rep in 1 2 3; do
  for S in 0 2; do TAHOE_CLUSTER=$S timeout 200 python tools/probe_prof.py C3 10000000 >> $L 2>&1; done
done
sed "s/checksum.*//" $L | cut -c1-200
TAHOE_CLUSTER=2 timeout 600 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed --clock-control none \
   -k regex:k_smem_forest -s 1 -c 1 --csv --log-file gpurun_out/r2h_ncu_cluster2.csv python tools/run_one.py C3 4 0 3 > gpurun_out/r2h_ncu_cluster2.log 2>&1; echo "ncu rc=$?"; tail -5 gpurun_out/r2h_ncu_cluster2.csv | cut -c1-300
(time timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r2h_bench.json 2> gpurun_out/r2h_bench.err); echo "bench rc=$?"; head -c 300 gpurun_out/r2h_bench.json; echo
(time timeout 600 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/r2h_bench_ref.json 2> gpurun_out/r2h_bench_ref.err); echo "ref rc=$?"; cat gpurun_out/r2h_bench_ref.json | cut -c1-700
