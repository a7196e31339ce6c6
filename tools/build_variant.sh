#!/bin/bash
# tools/build_variant.sh NAME "-DTAHOE_GROUP=8 -DTAHOE_WARPS=8 ..." : experimental build of the whole
# library with extra defines into tahoe_b200/lib/libtahoe_b200_NAME.so (select it with
# TAHOE_B200_LIB=...).  For tuning runs only.
set -e
cd "$(dirname "$0")/../tahoe_b200"
name=$1; defs=$2
mkdir -p build/$name lib
pids=()
for f in abi packer textio gen kernels_direct kernels_shared_data kernels_smem; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $defs -c csrc/$f.cu -o build/$name/$f.o &
  pids+=($!)
done
for p in "${pids[@]}"; do wait $p; done
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o lib/libtahoe_b200_$name.so build/$name/*.o -cudart static
echo lib/libtahoe_b200_$name.so
