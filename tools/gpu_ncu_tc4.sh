cd $GRAFT_REPO_ROOT
timeout 120 python tools/prof_tc.py 1024 3 > gpurun_out/prof_tc_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:bf_tc4_scan -s 1 -c 1 -o gpurun_out/prof_bf_tc4 python tools/prof_tc.py 1024 3 > gpurun_out/ncu_tc4.log 2>&1
cat gpurun_out/prof_tc_plain.log; tail -3 gpurun_out/ncu_tc4.log
