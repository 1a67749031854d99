#!/bin/bash
set -u
XP_GAE_KERNEL=spec timeout 150 python -m pytest tests/test_kernels_gpu.py tests/test_rollout_scenarios_gpu.py tests/test_components_gpu.py -m gpu -q -x -k "gae or rollout or Rollout or returns or arena" 2>&1 | tail -3
for P in 1 4 16 256; do
  S=$(( P < 16 ? 64 / P : (P == 16 ? 8 : 2) ))
  for K in tma spec; do echo -n "P=$P $K: "; XP_GAE_KERNEL=$K timeout 60 tools/micro/gae_bench0 $P $S | tail -2 | head -1; done
done
for K in tma spec; do echo "== graph-timed $K"; XP_GAE_KERNEL=$K timeout 200 python tools/micro/c3.py 2>&1 | tail -2; done
