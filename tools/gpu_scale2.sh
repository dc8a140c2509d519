# N = 1, 2 (and 4, 8 when the box has them) with the sharded C4 batch in the line
cd $GRAFT_REPO_ROOT
NG=$(nvidia-smi -L | wc -l)
[ "$SKIP_N1" = 1 ] || (timeout 1200 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "n1 exit $?")
for n in 2 4 8; do
  if [ $n -le $NG ]; then
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29540+n)) bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "n$n exit $?"
  fi
done
python - <<'P'
import json,glob
for f in sorted(glob.glob('gpurun_out/bench_n[1248].json')):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][-1])
        c=d['c4_local_window_batch']
        print(f, round(d['value']), round(d['e2e']['value']), 'c4 strong ms', round(c['strong']['ms_per_batch'],3), 'weak kf/s', round(c['weak']['keyframes_per_s']), c['strong']['matches_all_ranks'], c['weak']['matches_all_ranks'])
        if 'also' in d: print(d['also']['n4_bag_of_words']['score_l1_1000_keyframes'])
    except Exception as e: print(f,'failed',e)
P
