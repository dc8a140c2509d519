# GPU run 8: bench N=1 (default step = 32768 pairs), reference arm, ncu launch list of a short bench
cd $GRAFT_REPO_ROOT
timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?" >> gpurun_out/bench_n1.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
B="python bench.py --steps 2 --warmup 1 --pairs-per-step 8192 --no-also"
timeout 300 $B > gpurun_out/ncu_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu_launches.log 2>&1
tail -3 gpurun_out/bench_n1.err gpurun_out/ncu_launches.log; head -c 2500 gpurun_out/bench_n1.json
