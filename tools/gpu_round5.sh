#!/bin/bash
# end-of-round evidence: GPU tests, bench (both arms), launch list, full ncu captures. Outputs -> gpurun_out/
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py --steps 300 --warmup 10 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cut -c1-1200 gpurun_out/bench.json; tail -3 gpurun_out/bench.err
timeout 300 python bench.py --impl reference --steps 300 --warmup 10 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cut -c1-600 gpurun_out/bench_ref.json
SMALL="python bench.py --steps 3 --warmup 3 --rows 131072 --no-also --no-cpu --no-e2e"
timeout 600 $SMALL > gpurun_out/small_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'move_kernel|per_sample|set_|gae_|tree_|randperm|philox' -c 400 --csv --log-file gpurun_out/launches_bench.csv $SMALL > gpurun_out/ncu_small.log 2>&1
echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'move_kernel<\(int\)0>|per_sample|set_mid' -s 12 -c 8 -f -o gpurun_out/prof_c2 $SMALL > gpurun_out/ncu_c2.log 2>&1
echo "ncu c2 rc=$?"
Q="python tools/quick_bench.py per24 gae"
timeout 600 $Q > gpurun_out/quick2.log 2>&1 && {
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'gae_tma' -s 10 -c 2 -f -o gpurun_out/prof_gae $Q > gpurun_out/ncu_gae.log 2>&1; echo "ncu gae rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'per_sample_kernel' -s 10 -c 2 -f -o gpurun_out/prof_sample $Q > gpurun_out/ncu_sample.log 2>&1; echo "ncu sample rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'set_hist|set_scan|set_scatter|set_wbucket|set_dense' -s 24 -c 12 -f -o gpurun_out/prof_update $Q > gpurun_out/ncu_update.log 2>&1; echo "ncu update rc=$?"
}
cat gpurun_out/quick2.log
G5="python tools/micro/gather_c5.py"
timeout 300 $G5 > gpurun_out/gather_c5.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:'move_kernel' -s 6 -c 2 -f -o gpurun_out/prof_gather_c5 $G5 > gpurun_out/ncu_gather_c5.log 2>&1; echo "ncu c5 rc=$?"
cat gpurun_out/gather_c5.log
