cd $GRAFT_REPO_ROOT
timeout 400 python tools/tc_check.py --variant 0 > gpurun_out/tc_check4.log 2>&1
for v in 0 16; do echo "variant $v"; SLAMGPU_TC_VARIANT=$v timeout 120 python tools/prof_tc.py 2048 4; done > gpurun_out/tc_exp.log 2>&1
timeout 300 python tools/tc_sizes.py > gpurun_out/tc_sizes.log 2>&1
grep -v "mismatching rows 0" gpurun_out/tc_check4.log | tail -20; cat gpurun_out/tc_exp.log gpurun_out/tc_sizes.log
