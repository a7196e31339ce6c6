#!/bin/bash
# tools/build_deep_variant.sh NAME "-DTAHOE_DEEP_WALKS_512=2 ..." : tuning build of k_smem_forest_deep
# -> tahoe_b200/lib/libtahoe_b200_NAME.so (select it with TAHOE_B200_LIB=...).  Needs the normal build
# first (python -m tahoe_b200.build): only the per-depth objects are recompiled.
set -e
cd "$(dirname "$0")/../tahoe_b200"
N=$1; DEFS=$2; mkdir -p build/$N
F="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-O2,-fno-fast-math --fmad=true -prec-div=true -prec-sqrt=true -ftz=false"
pids=()
for d in $(seq 1 14); do
  nvcc $F -DTAHOE_DEPTH=$d $DEFS -c csrc/kernels_smem_depth.cu -o build/$N/kernels_smem_depth_$d.o &
  pids+=($!)
done
for p in "${pids[@]}"; do wait $p; done
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o lib/libtahoe_b200_$N.so build/*.cu.o build/$N/*.o -cudart static
echo lib/libtahoe_b200_$N.so
