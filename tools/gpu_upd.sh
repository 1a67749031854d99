#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "tree or per or Per or prior" 2>&1 | tail -3
timeout 300 python tools/micro/c4.py 2>&1 | tail -1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'set_' -s 20 -c 5 --csv --log-file gpurun_out/launches_update.csv python tools/quick_bench.py per24 > /dev/null 2>&1; grep -o 'set_[a-z]*_kernel[^"]*"\|"[0-9.]*"$' gpurun_out/launches_update.csv | paste - - | head -12
