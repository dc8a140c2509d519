#!/bin/bash
# ncu --set full of the FP4 scan kernel on a short C5-shaped batch (the plain run first, as the recipe asks)
t=${1:-1}
SLAMGPU_TC4_TEAMS=$t timeout 200 python tools/prof_tc.py 1024 3 > gpurun_out/ncu_plain_$t.log 2>&1 &&
SLAMGPU_TC4_TEAMS=$t timeout 600 ncu --set full --clock-control none --import-source on -k regex:bf_tc4_scan -s 2 -c 1 -f -o gpurun_out/tc4_teams$t python tools/prof_tc.py 1024 3 > gpurun_out/ncu_run_$t.log 2>&1
tail -2 gpurun_out/ncu_plain_$t.log gpurun_out/ncu_run_$t.log
