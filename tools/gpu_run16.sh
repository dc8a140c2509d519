cd $GRAFT_REPO_ROOT
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
timeout 600 python -m pytest tests/test_gpu_voc.py -m gpu -q -x 2>&1 | tail -2
timeout 1200 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"; tail -n 3 gpurun_out/bench_n1.err
python -c "
import json; d=json.load(open('gpurun_out/bench_n1.json')); print(d['value'], d['e2e']['value'], d['c4_local_window_batch']); print(d['also']['n4_bag_of_words']['score_l1_1000_keyframes']); print(d['also']['n4_bag_of_words']['bow_of_8000_features'])"
