#!/bin/bash
# One GPU call: tests, bench (both arms), launch list + full ncu captures. Outputs -> gpurun_out/
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 200 --warmup 10 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 200 --warmup 10 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cat gpurun_out/bench_ref.json
timeout 600 python tools/quick_bench.py > gpurun_out/quick.log 2>&1; echo "quick rc=$?"; cat gpurun_out/quick.log
SMALL="python bench.py --steps 3 --warmup 3 --rows 131072 --no-also --no-cpu --no-e2e"
timeout 600 $SMALL > gpurun_out/small_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'move_kernel|per_sample|set_|gae_|tree_|randperm|philox' -c 400 --csv --log-file gpurun_out/launches_bench.csv $SMALL > gpurun_out/ncu_small.log 2>&1
echo "ncu launches rc=$?"
timeout 600 $SMALL > gpurun_out/small_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'move_kernel' -s 8 -c 2 -o gpurun_out/prof_gather $SMALL > gpurun_out/ncu_gather.log 2>&1
echo "ncu gather rc=$?"
timeout 600 python tools/quick_bench.py per gae > gpurun_out/quick2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'per_sample|set_|gae_' -s 40 -c 12 -o gpurun_out/prof_tree_gae python tools/quick_bench.py per gae > gpurun_out/ncu_tree.log 2>&1
echo "ncu tree rc=$?"
ls -la gpurun_out
