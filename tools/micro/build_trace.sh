#!/bin/bash
# developer build: libb200xp with gae.cu compiled -DXP_GAE_TRACE (per-CTA phase stamps), into tools/micro/_trace/
set -eu
cd "$(dirname "$0")/../.."
python -m agilerl_b200.build > /dev/null
OUT=tools/micro/_trace
mkdir -p $OUT
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC,-O2 \
  -I include -I agilerl_b200/csrc -DXP_BUILDING=1 ${XP_TRACE_ON--DXP_GAE_TRACE} ${XP_TRACE_DEFS:-} -c agilerl_b200/csrc/gae.cu -o $OUT/gae_trace.o
OBJS=$(ls agilerl_b200/_lib/obj/*.o | grep -v "/gae.o")
nvcc -shared -o $OUT/libb200xp_trace.so $OBJS $OUT/gae_trace.o -gencode arch=compute_100a,code=sm_100a -cudart static
echo built $OUT/libb200xp_trace.so
