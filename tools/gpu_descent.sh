#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/team_test.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/team_test.log
timeout 300 python tools/micro/c4.py 2>&1 | tail -1
