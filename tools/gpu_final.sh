#!/bin/bash
# end-of-round check as the driver runs it: GPU tests, smoke, bench (both arms). Outputs -> gpurun_out/
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
python -c "import json; d=json.loads(open('gpurun_out/bench_ref.json').read()); print('reference', round(d['value']), d['cpu_baseline']['kind'], 'wall', d['wall_s'], '| upper', round(d['population_parallel_upper_bound']['value']), d['population_parallel_upper_bound']['cores'])"
timeout 1200 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python tools/show_bench.py gpurun_out/bench.json
# capture of the dominant C3 kernel (gae_tma_kernel, 16 agents) with the final code
G="python tools/micro/gae_once.py 16"
timeout 120 $G > gpurun_out/gae_once.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:gae_tma -s 2 -c 2 -f -o gpurun_out/prof_gae $G > gpurun_out/ncu_gae.log 2>&1; echo "ncu gae rc=$?"
