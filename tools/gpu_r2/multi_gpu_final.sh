#!/bin/bash
# N GPUs on the final library: the fused-gather test, then the sharded bench (no per-config sub-records at N > 1)
mkdir -p gpurun_out
N=${1:-2}
timeout 600 python -m pytest tests/test_peer_gather_gpu.py -x -q > gpurun_out/r2f_n${N}_pytest.log 2>&1; echo "pytest(peer gather) rc=$?"; tail -2 gpurun_out/r2f_n${N}_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2f_bench_C3_n$N.json 2> gpurun_out/r2f_bench_n$N.err; echo "bench N=$N rc=$?"; grep -v "^\[W\|^W1105" gpurun_out/r2f_bench_n$N.err | tail -c 400; cut -c1-700 gpurun_out/r2f_bench_C3_n$N.json
