#!/bin/bash
# tools/build_variant.sh NAME "-DTAHOE_GROUP=8 ..." : experimental build of kernels_smem.cu with extra
# defines, linked with the other objects into tahoe_b200/lib/libtahoe_b200_NAME.so (select it with
# TAHOE_B200_LIB=...).  For tuning runs only.
set -e
cd "$(dirname "$0")/../tahoe_b200"
name=$1; defs=$2
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC $defs -c csrc/kernels_smem.cu -o build/kernels_smem_$name.o
objs=$(ls build/*.cu.o | grep -v kernels_smem.cu.o)
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o lib/libtahoe_b200_$name.so $objs build/kernels_smem_$name.o -cudart static
echo lib/libtahoe_b200_$name.so
