#!/bin/bash
set -u
for k in 0 1 2; do echo "== XP_GAE_PDL=$k"; XP_GAE_PDL=$k timeout 300 python tools/micro/c3.py 2>&1 | tail -3; done
XP_GAE_PDL=2 timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_rollout_scenarios_gpu.py -m gpu -q -k "gae or rollout or arena" 2>&1 | tail -2
echo "== c4 (single dense launch)"; timeout 300 python tools/micro/c4.py 2>&1 | tail -1
timeout 600 python -m pytest tests/test_kernels_gpu.py -m gpu -x -q -k "tree or per" 2>&1 | tail -2
