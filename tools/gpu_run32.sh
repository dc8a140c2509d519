#!/bin/bash
# N3 (PnP verification support) GPU validation after the container was re-created: its parity tests, the whole GPU
# suite, smoke, the bench line, launch time and one full ncu capture of the scoring kernel.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
timeout 400 python -m pytest tests/test_gpu_pnp.py -x -q -m gpu 2>&1 | tail -15
timeout 200 python tools/prof_pnp.py 5 2>&1 | tail -3
timeout 200 python __graft_entry__.py smoke 2>&1 | tail -3
timeout 500 python bench.py > gpurun_out/bench_n1_run32.json 2> gpurun_out/bench_n1_run32.err; echo "bench rc=$?"; tail -c 2500 gpurun_out/bench_n1_run32.json; tail -3 gpurun_out/bench_n1_run32.err
timeout 300 ncu --set full --clock-control none --import-source on -k regex:pnp_score_kernel -c 1 -o gpurun_out/pnp_score_full -f python tools/prof_pnp.py 1 > gpurun_out/ncu_pnp.log 2>&1; echo "ncu rc=$?"
(time timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -5) 2>&1
