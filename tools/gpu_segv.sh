#!/bin/bash
set -u
mkdir -p gpurun_out
G="tests/test_kernels_gpu.py tests/test_components_gpu.py tests/test_hostaccel_gpu.py tests/test_marb_gpu.py tests/test_peer_gpu.py"
run() { name=$1; shift; SEGV_BT=1 LD_PRELOAD=$PWD/tools/dbg/segv_bt.so timeout 600 python -m pytest "$@" -m gpu -x -q > gpurun_out/segv_$name.log 2>&1; echo "$name rc=$?"; grep -v "^\.\|^$" gpurun_out/segv_$name.log | tail -50; }
run full tests
run abi tests/test_abi_cpu.py $G
run host tests/test_host_cpu.py $G
run og tests/test_oracle_golden.py $G
run ovr tests/test_oracle_vs_reference_cpu.py $G
run gloo tests/test_population_gloo_cpu.py $G
