#!/bin/bash
# bench.py on N GPUs of one box, launched as the driver does.  usage: tools/gpu_scale.sh N [extra bench args]
set -u
N=${1:-2}; shift || true
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 200 --warmup 10 "$@" > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "bench n=$N rc=$?"; grep -v "^W1\|^\[W\|NCCL\|^$" gpurun_out/bench_n$N.err | tail -8
python tools/show_bench.py gpurun_out/bench_n$N.json
