#!/bin/bash
# round 2, call A: parity of the new tree-set / dense / lanes-GAE code + micro-benchmarks. Outputs -> gpurun_out/
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv,noheader
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
XP_GAE_KERNEL=lanes timeout 600 python -m pytest tests/test_kernels_gpu.py tests/test_rollout_scenarios_gpu.py tests/test_components_gpu.py -m gpu -q -k "gae or rollout or Rollout or returns" > gpurun_out/pytest_lanes.log 2>&1; echo "pytest lanes rc=$?"; tail -3 gpurun_out/pytest_lanes.log
echo "== gae tma"; XP_GAE_KERNEL=tma timeout 120 tools/micro/gae_bench0 16 8 | tail -2
echo "== gae lanes"; XP_GAE_KERNEL=lanes timeout 120 tools/micro/gae_bench0 16 8 | tail -2
echo "== gae lanes x16"; XP_GAE_KERNEL=lanes timeout 120 tools/micro/gae_bench0 256 2 | tail -1
echo "== gae tma x16"; XP_GAE_KERNEL=tma timeout 120 tools/micro/gae_bench0 256 2 | tail -1
echo "== gae lanes P=1"; XP_GAE_KERNEL=lanes timeout 120 tools/micro/gae_bench0 1 64 | tail -1
echo "== gae tma P=1"; XP_GAE_KERNEL=tma timeout 120 tools/micro/gae_bench0 1 64 | tail -1
timeout 60 tools/micro/stream_floor 16 8 | tail -3
echo "== c3 graph-timed (bench path)"; for k in tma lanes; do XP_GAE_KERNEL=$k timeout 300 python tools/micro/c3.py 2>&1 | tail -3; done
echo "== c4 graph-timed"; for d in 0 -1 1; do echo "top_delta $d"; XP_TOP_DELTA=$d timeout 300 python tools/micro/c4.py 2>&1 | tail -1; done
echo "== update sizes (cap 2^20)"; timeout 300 python tools/micro/update_sizes.py 2>&1 | tail -6
echo "== skew"; timeout 300 python tools/micro/update_skew.py 2>&1 | tail -6
