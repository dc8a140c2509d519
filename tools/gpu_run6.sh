# GPU run 6: full GPU suite, bench N=1, reference arm, ncu launch list + full captures (scan v2, proj v2)
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?" >> gpurun_out/bench_n1.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
B="python bench.py --steps 2 --warmup 1 --pairs-per-step 512 --no-also"
timeout 300 $B > gpurun_out/ncu_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_launches.log 2>&1
timeout 300 $B > gpurun_out/ncu_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:bf_scan -s 2 -c 2 -o gpurun_out/prof_bf_scan_v2 $B > gpurun_out/ncu_full.log 2>&1
timeout 120 python tools/prof_proj.py 8 > gpurun_out/prof_proj_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:proj_match -s 4 -c 2 -o gpurun_out/prof_proj_match_v2 python tools/prof_proj.py 8 > gpurun_out/ncu_proj.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; tail -3 gpurun_out/bench_n1.err gpurun_out/ncu_launches.log gpurun_out/ncu_full.log gpurun_out/ncu_proj.log
