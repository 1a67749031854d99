#!/bin/bash
# round 2, call C: full GPU suite on the reworked tree set + C4 / skew / size micro-benchmarks
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
echo "== c4"; timeout 300 python tools/micro/c4.py 2>&1 | tail -1
echo "== skew"; timeout 300 python tools/micro/update_skew.py 2>&1 | tail -6
echo "== sizes"; timeout 300 python tools/micro/update_sizes.py 2>&1 | tail -6
