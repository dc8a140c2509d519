#!/bin/bash
# The whole C5 job (499,500 pairs, resident database, pinned result arrays) strong-scaled over the GPUs of the box.
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
NG=$(nvidia-smi -L | wc -l)
: > gpurun_out/c5_full_job_scale.json
for n in ${NS:-1 2 4 8}; do
  if [ $n -le $NG ]; then
    if [ $n = 1 ]; then timeout 300 python tools/c5_full_job.py 8192 pinned resident 2>&1 | tail -1 | tee -a gpurun_out/c5_full_job_scale.json
    else timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) tools/c5_full_job.py 8192 pinned resident 2>&1 | grep '^{' | tee -a gpurun_out/c5_full_job_scale.json
    fi
  fi
done
