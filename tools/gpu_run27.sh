#!/bin/bash
# N=2 bench with the NCCL database broadcast (launch as the driver does)
cd $GRAFT_REPO_ROOT
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r27_bench_n2.json 2> gpurun_out/r27_bench_n2.err; echo "bench rc=$?"
tail -c 600 gpurun_out/r27_bench_n2.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r27_bench_n2.json').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ('value','n_gpus','ms_per_step','e2e','replicate','clocks')})
print(d.get('c4_local_window_batch'))
P
