#!/bin/bash
# N GPUs: the fused gather (kernel epilogue stores into rank 0's IPC-mapped buffer) test, then the sharded bench
mkdir -p gpurun_out
N=${1:-2}
nvidia-smi -L | head -8
timeout 600 python -m pytest tests/test_peer_gather_gpu.py tests/test_shard_gloo.py -x -q > gpurun_out/r2_n2_pytest.log 2>&1; echo "pytest(peer gather) rc=$?"; tail -4 gpurun_out/r2_n2_pytest.log
if [ "$N" = "2" ]; then
  timeout 900 python -m pytest tests/test_importers.py tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/r2_n2_pytest_imp.log 2>&1; echo "pytest(importers, parity) rc=$?"; tail -4 gpurun_out/r2_n2_pytest_imp.log
fi
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2_bench_C3_n$N.json 2> gpurun_out/r2_bench_n$N.err; echo "bench N=$N rc=$?"; grep -v "^\[W\|^W1105" gpurun_out/r2_bench_n$N.err | tail -c 600; cat gpurun_out/r2_bench_C3_n$N.json | cut -c1-1800
