#!/bin/bash
run() { env TAHOE_SPARSE_CUT=$1 TAHOE_SPARSE_TPC=$2 python tools/debug_sparse.py $3 0 ${4:-64} 2>&1 | tail -1; }
run 32 1 75776
run 256 1 75781
run 32 3 151552
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -k "sparse" 2>&1 | tail -3
