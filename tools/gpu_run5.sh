cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?" >> gpurun_out/bench_n1.err
timeout 120 python tools/prof_proj.py 8 > gpurun_out/prof_proj_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:proj_match -s 4 -c 2 -o gpurun_out/prof_proj_match python tools/prof_proj.py 8 > gpurun_out/ncu_proj.log 2>&1
tail -n 4 gpurun_out/pytest_gpu.log; tail -n 3 gpurun_out/bench_n1.err gpurun_out/ncu_proj.log
