#!/bin/bash
set -u
mkdir -p gpurun_out
Q="python tools/quick_bench.py per20 per24 gae"
timeout 600 $Q > gpurun_out/quick2.log 2>&1 && {
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'gae_pipe' -s 10 -c 1 -o gpurun_out/prof_gae2 $Q > gpurun_out/ncu_gae.log 2>&1; echo "ncu gae rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'set_small' -s 10 -c 1 -o gpurun_out/prof_small2 $Q > gpurun_out/ncu_small.log 2>&1; echo "ncu small rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'set_ticket|set_scatter|set_bucket|set_dense' -s 12 -c 4 -o gpurun_out/prof_large2 $Q > gpurun_out/ncu_large.log 2>&1; echo "ncu large rc=$?"
}
cat gpurun_out/quick2.log
