#!/bin/bash
# bench.py at N GPUs (arg 1), launched the way the driver does; output -> gpurun_out/bench_nN.json
set -u
N=$1
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 200 --warmup 10 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench N=$N rc=$?"
grep -v "^\*\|^$\|NCCL version" gpurun_out/bench_n$N.json | cut -c1-1500; tail -3 gpurun_out/bench_n$N.err
