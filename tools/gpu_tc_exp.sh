cd $GRAFT_REPO_ROOT
timeout 400 python tools/tc_check.py --variant 0 > gpurun_out/tc_check4.log 2>&1
for v in 0 32 36 8; do echo "variant $v"; SLAMGPU_TC_VARIANT=$v timeout 120 python tools/prof_tc.py 2048 4; done > gpurun_out/tc_exp.log 2>&1
grep -v "mismatching rows 0" gpurun_out/tc_check4.log | tail -12; cat gpurun_out/tc_exp.log
