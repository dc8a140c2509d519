#!/bin/bash
# Resident key-frame database (slamgpu_kfdb_*): parity tests, then the whole C5 job with and without it.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
(time timeout 600 python -m pytest tests/test_gpu_bf.py -x -q -m gpu 2>&1 | tail -8) 2>&1
timeout 600 python tools/c5_full_job.py 8192 pinned 2>&1 | tail -2 | tee gpurun_out/c5_full_job.json
timeout 600 python tools/c5_full_job.py 8192 pinned resident 2>&1 | tail -2 | tee -a gpurun_out/c5_full_job.json
timeout 600 python tools/c5_full_job.py 32768 pinned resident 2>&1 | tail -1 | tee -a gpurun_out/c5_full_job.json
timeout 600 python tools/c5_full_job.py 2048 pinned resident 2>&1 | tail -1 | tee -a gpurun_out/c5_full_job.json
