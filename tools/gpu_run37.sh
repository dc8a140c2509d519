#!/bin/bash
# Full GPU suite, smoke, default bench line (with the resident-database leg) and the reference arm.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(time timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -5) 2>&1
timeout 200 python __graft_entry__.py smoke 2>&1 | tail -3
(time timeout 600 python bench.py > gpurun_out/bench_n1_run37.json 2> gpurun_out/bench_n1_run37.err) 2>&1 | tail -3; echo "bench rc=$?"; head -c 1800 gpurun_out/bench_n1_run37.json; tail -3 gpurun_out/bench_n1_run37.err
timeout 300 python bench.py --impl reference > gpurun_out/bench_ref_n1_run37.json 2>/dev/null; head -c 600 gpurun_out/bench_ref_n1_run37.json
timeout 600 python tools/c5_full_job.py 8192 pinned resident 2>&1 | tail -1 | tee gpurun_out/c5_full_job_v2.json
