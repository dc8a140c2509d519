# 1 -> 8 GPU weak-scaling run of bench.py (launch exactly as the driver does)
cd $GRAFT_REPO_ROOT
for n in 4 8; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29520+n)) bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "exit $?" >> gpurun_out/bench_n$n.err
head -c 200 gpurun_out/bench_n$n.json; echo
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 bench.py --impl reference --gpus 8 --steps 2 --warmup 1 > gpurun_out/bench_ref_n8.json 2> gpurun_out/bench_ref_n8.err; echo "exit $?" >> gpurun_out/bench_ref_n8.err
tail -n 2 gpurun_out/bench_n4.err gpurun_out/bench_n8.err gpurun_out/bench_ref_n8.err
