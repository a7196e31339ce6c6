#!/bin/bash
# round 2, call Q: the peer-gather test on ONE GPU (two processes, same device), then the whole suite as the driver runs it
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_peer_gather_gpu.py -x -q > gpurun_out/r2q_peer.log 2>&1; echo "peer gather (one GPU) rc=$?"; tail -5 gpurun_out/r2q_peer.log | cut -c1-200
timeout 1200 python -m pytest tests/ -x -q -m gpu > gpurun_out/r2q_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2q_pytest.log
