#!/bin/bash
# final check of the round: GPU tests, smoke, bench (both arms), launch list. Outputs -> gpurun_out/
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py --steps 300 --warmup 10 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cut -c1-400 gpurun_out/bench.json; tail -3 gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 300 --warmup 10 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cut -c1-200 gpurun_out/bench_ref.json
SMALL="python bench.py --steps 3 --warmup 3 --rows 131072 --no-also --no-cpu --no-e2e"
timeout 600 $SMALL > gpurun_out/small_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'move_kernel|per_sample|set_|gae_|tree_|randperm|philox' -c 400 --csv --log-file gpurun_out/launches_bench.csv $SMALL > gpurun_out/ncu_small.log 2>&1
echo "ncu launches rc=$?"
