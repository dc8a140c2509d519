#!/bin/bash
# Final validation of the tree: full GPU suite, smoke, loop-query latency against the resident database, bench line.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(time timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -5) 2>&1
timeout 200 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 300 python tools/time_kfdb_query.py 2>&1 | tail -5 | tee gpurun_out/kfdb_query_latency.txt
timeout 600 python bench.py > gpurun_out/bench_n1_run40.json 2> gpurun_out/bench_n1_run40.err; echo "bench rc=$?"; head -c 1500 gpurun_out/bench_n1_run40.json | tail -c 700
