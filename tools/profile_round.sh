#!/bin/bash
# The profiling evidence of a round, on one B200 (each ncu pass only after the same command ran clean without ncu):
#   1. launch list of the bench command (share of the step per kernel)
#   2. ncu --set full of the dominant kernel (FP4 scan) on a C5-shaped batch
#   3. ncu --set full of the projection search on a 200k-point map
tag=${1:-r02}
python bench.py --no-also --steps 2 --warmup 3 --pairs-per-step 8192 > gpurun_out/${tag}_plain_bench.log 2> /dev/null &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${tag}_ncu_launches_bench.csv \
    python bench.py --no-also --steps 2 --warmup 3 --pairs-per-step 8192 > gpurun_out/${tag}_ncu_bench.log 2>&1
python tools/prof_tc.py 1024 3 > gpurun_out/${tag}_plain_tc.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:bf_tc4_scan -s 2 -c 1 -f -o gpurun_out/${tag}_bf_tc4_scan \
    python tools/prof_tc.py 1024 3 > gpurun_out/${tag}_ncu_tc.log 2>&1
python tools/prof_proj_big.py > gpurun_out/${tag}_plain_proj.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:proj_search -s 2 -c 1 -f -o gpurun_out/${tag}_proj_search \
    python tools/prof_proj_big.py > gpurun_out/${tag}_ncu_proj.log 2>&1
ls -la gpurun_out/${tag}_*
