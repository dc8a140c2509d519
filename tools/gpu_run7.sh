# GPU run 7: full GPU suite (both brute-force engines), bench N=1 with the tensor-core scan, ncu of the CTA-pair kernel
cd $GRAFT_REPO_ROOT
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?" >> gpurun_out/bench_n1.err
timeout 120 python tools/prof_tc.py 1024 3 > gpurun_out/prof_tc_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:bf_tc4_scan -s 1 -c 1 -o gpurun_out/prof_bf_tc4_final python tools/prof_tc.py 1024 3 > gpurun_out/ncu_tc.log 2>&1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 300 python tools/tc_sizes.py > gpurun_out/tc_sizes.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -3 gpurun_out/bench_n1.err; cat gpurun_out/prof_tc_plain.log; tail -3 gpurun_out/ncu_tc.log; head -c 1500 gpurun_out/bench_n1.json
