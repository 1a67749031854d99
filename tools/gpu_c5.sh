#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_components_gpu.py -m gpu -x -q -k "randperm or Replay or replay or uniform or sample" 2>&1 | tail -3
timeout 900 python bench.py --steps 50 --warmup 5 --rows 131072 --no-cpu --no-e2e > gpurun_out/bench_c5.json 2> gpurun_out/bench_c5.err; echo "rc=$?"; tail -3 gpurun_out/bench_c5.err
python -c "
import json; d=json.load(open('gpurun_out/bench_c5.json'))
print(json.dumps(d['also'][-1])[:900])
"
