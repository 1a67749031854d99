#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_n2.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu_n2.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/micro/peer_time.py > gpurun_out/peer_time_n2.log 2>&1; echo "peer_time rc=$?"; tail -4 gpurun_out/peer_time_n2.log
