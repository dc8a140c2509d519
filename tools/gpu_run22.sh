cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -n 3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_n1.json 2> gpurun_out/bench_ref_n1.err; echo "ref exit $?"
timeout 1200 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"
python tools/prof_voc.py && ncu --set full --clock-control none --import-source on -k regex:voc_transform -s 2 -c 1 -o gpurun_out/voc_transform python tools/prof_voc.py > gpurun_out/ncu_voc.log 2>&1
ncu -i gpurun_out/voc_transform.ncu-rep --page details > gpurun_out/voc_transform_details.txt 2>&1
grep -E "Duration|DRAM Throughput|L2 Cache Throughput|L1/TEX Cache Throughput|Issue Slots Busy|Registers Per|Achieved Occupancy|L2 Hit|Executed Ipc Active" gpurun_out/voc_transform_details.txt
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 2 --warmup 3 --no-also > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches exit $?"
