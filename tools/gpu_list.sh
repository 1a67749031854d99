#!/bin/bash
# launch list (ncu, per-kernel durations) of a quick_bench selection: tools/gpu_list.sh per24
mkdir -p gpurun_out
python tools/quick_bench.py "$@" > gpurun_out/q.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -k regex:"set_|per_sample|gae_|move_" -c 400 --csv --log-file gpurun_out/launch_list.csv python tools/quick_bench.py "$@" > /dev/null 2>&1
echo "ncu rc=$?"
