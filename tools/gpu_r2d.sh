#!/bin/bash
# round 2, call D: full GPU suite, then the round's ncu evidence (launch lists + full captures). Outputs -> gpurun_out/
set -u
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
C="python tools/micro/c4_once.py 4"
timeout 300 $C > gpurun_out/c4_once.log 2>&1 && {
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'set_|per_sample|move_kernel|gae_' --csv --log-file gpurun_out/launches_c4.csv $C > gpurun_out/ncu_l4.log 2>&1; echo "ncu launches c4 rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'set_slot|set_wbucket|set_dense' -s 9 -c 8 -f -o gpurun_out/prof_update $C > gpurun_out/ncu_update.log 2>&1; echo "ncu update rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'per_sample_kernel' -s 2 -c 2 -f -o gpurun_out/prof_sample $C > gpurun_out/ncu_sample.log 2>&1; echo "ncu sample rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'move_kernel' -s 1 -c 2 -f -o gpurun_out/prof_insert $C > gpurun_out/ncu_insert.log 2>&1; echo "ncu insert rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'gae_tma' -s 3 -c 2 -f -o gpurun_out/prof_gae $C > gpurun_out/ncu_gae.log 2>&1; echo "ncu gae rc=$?"
}
SMALL="python bench.py --steps 4 --warmup 3 --rows 131072 --no-also --no-cpu --no-e2e"
timeout 600 $SMALL > gpurun_out/small_plain.log 2>&1 && {
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'move_kernel|per_sample|set_|gae_|tree_|randperm|philox' -c 400 --csv --log-file gpurun_out/launches_bench.csv $SMALL > gpurun_out/ncu_small.log 2>&1; echo "ncu launches bench rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'move_kernel' -s 17 -c 2 -f -o gpurun_out/prof_gather $SMALL > gpurun_out/ncu_gather.log 2>&1; echo "ncu gather rc=$?"
}
G5="python tools/micro/gather_c5.py"
timeout 300 $G5 > gpurun_out/gather_c5.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:'move_kernel' -s 6 -c 2 -f -o gpurun_out/prof_gather_c5 $G5 > gpurun_out/ncu_gather_c5.log 2>&1; echo "ncu c5 rc=$?"
ls -la gpurun_out/*.ncu-rep
