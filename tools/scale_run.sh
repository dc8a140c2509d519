#!/bin/bash
# bench.py at N GPUs exactly as the driver launches it; line -> gpurun_out/bench_n$N.json
n=$1; shift
if [ "$n" = "1" ]; then python bench.py --gpus 1 --no-also "$@" > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err
else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n "$@" > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; fi
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_n$n.json").read().strip().splitlines()[-1])
print("N=$n value %.0f e2e %.0f res %.0f job %.4f s (%d matches) replicate %s c4 %s" % (d["value"], d["e2e"]["value"], d["e2e_resident_database"]["value"], d["c5_whole_job_strong"]["seconds"], d["c5_whole_job_strong"]["matches_total"], d.get("replicate"), {k:(v.get("ms_per_batch") if isinstance(v,dict) else v) for k,v in d["c4_local_window_batch"].items()}))
PY
