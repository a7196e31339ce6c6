#!/bin/bash
# k_smem_forest_deep: plan x top-depth sweep on config 4 (PAIR_PLANS, PAIR_DEPTHS, PAIR_LIBS = variant libraries)
mkdir -p gpurun_out; L=gpurun_out/pair.log; : > $L
for lib in ${PAIR_LIBS:-default}; do
  echo "== lib $lib" >> $L
  [ "$lib" = default ] && unset TAHOE_B200_LIB || export TAHOE_B200_LIB=tahoe_b200/lib/libtahoe_b200_$lib.so
  for plan in ${PAIR_PLANS:-"16,1,6" "16,1,4" "8,1,6"}; do
    for d in ${PAIR_DEPTHS:-6 7 8}; do
      TAHOE_SMEM_PLAN=$plan TAHOE_QTOP_DEPTH=$d timeout 300 python tools/probe_qtop.py 1000000 >> $L 2>&1
    done
  done
done
cut -c1-40,60-70,100-400 $L
