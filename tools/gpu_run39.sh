#!/bin/bash
# Launch list of the current bench (resident-database leg included) for profiles/.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-also > gpurun_out/bench_noalso.json 2> gpurun_out/bench_noalso.err; echo "bench rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_bench_run39.csv python bench.py --steps 2 --warmup 3 --no-also > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches exit $?"
python tools/summarise_launches.py gpurun_out/launches_bench_run39.csv | tee gpurun_out/launches_summary_run39.txt | head -20
