cd $GRAFT_REPO_ROOT
for v in 8 264; do echo "variant $v (N=160)"; SLAMGPU_TC_VARIANT=$v timeout 120 python tools/prof_tc.py 2048 4; done > gpurun_out/tc_exp.log 2>&1
cat gpurun_out/tc_exp.log
