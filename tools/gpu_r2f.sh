#!/bin/bash
set -u
for rep in 1 2; do for v in "" _g64x2 _g128x1 _g128x2; do
  echo "== libb200xp$v"
  B200XP_LIB=$PWD/agilerl_b200/_lib/libb200xp$v.so timeout 300 python tools/micro/c3.py 2>&1 | tail -2
done; done
