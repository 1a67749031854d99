#!/bin/bash
set -u
mkdir -p gpurun_out
echo "== previous library (HEAD tree_set.cuh) on the varying-batch test"
B200XP_LIB=$PWD/tools/dbg/libb200xp_prev.so XP_NO_HOSTACCEL=1 timeout 300 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "sizes_may_change" 2>&1 | tail -4
echo "== current library"
timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_components_gpu.py tests/test_hostaccel_gpu.py -m gpu -x -q -k "tree or per or Per or prior" 2>&1 | tail -3
timeout 300 python tools/micro/c4.py 2>&1 | tail -1
