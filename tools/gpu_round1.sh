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
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'move_kernel' -s 14 -c 2 -o gpurun_out/prof_gather $SMALL > gpurun_out/ncu_gather.log 2>&1
echo "ncu gather rc=$?"
Q="python tools/quick_bench.py per24 gae"
timeout 600 $Q > gpurun_out/quick2.log 2>&1 && {
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'gae_staged' -s 10 -c 2 -o gpurun_out/prof_gae $Q > gpurun_out/ncu_gae.log 2>&1; echo "ncu gae rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'per_sample_kernel' -s 10 -c 2 -o gpurun_out/prof_sample $Q > gpurun_out/ncu_sample.log 2>&1; echo "ncu sample rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'set_' -s 15 -c 6 -o gpurun_out/prof_update $Q > gpurun_out/ncu_update.log 2>&1; echo "ncu update rc=$?"
}
ls -la gpurun_out
