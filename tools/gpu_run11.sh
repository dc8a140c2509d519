# full GPU suite + bench + full C3 sequence + variant timing + ncu capture of the final K4 on the 200k map
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -n 3 gpurun_out/pytest_gpu.log
timeout 600 python tools/time_proj.py > gpurun_out/proj_variants_final.txt 2>&1; grep default gpurun_out/proj_variants_final.txt
timeout 900 python tools/bench_c3.py --frames 2000 > gpurun_out/c3_full.json 2> gpurun_out/c3_full.err; echo "c3 exit $?"; head -c 900 gpurun_out/c3_full.json; echo
timeout 900 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"; tail -n 3 gpurun_out/bench_n1.err
python tools/prof_proj_big.py && ncu --set full --clock-control none --import-source on -k regex:proj_match_group -s 3 -c 1 -o gpurun_out/proj_big_final python tools/prof_proj_big.py > gpurun_out/ncu_proj_big.log 2>&1
ncu -i gpurun_out/proj_big_final.ncu-rep --page details > gpurun_out/proj_big_final_details.txt 2>&1
