#!/bin/bash
# two GPUs: the fused gather (kernel epilogue stores into rank 0's IPC-mapped buffer) test, then the sharded bench
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m pytest tests/test_peer_gather_gpu.py tests/test_shard_gloo.py -x -q > gpurun_out/r2_n2_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_n2_pytest.log
N=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2_bench_C3_n$N.json 2> gpurun_out/r2_bench_n$N.err; echo "bench N=$N rc=$?"; tail -c 400 gpurun_out/r2_bench_n$N.err; cat gpurun_out/r2_bench_C3_n$N.json | cut -c1-1500
