# full GPU suite + C3 sequence (KD threads 1 vs 4) + bench
cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -n 3 gpurun_out/pytest_gpu.log
SLAMGPU_KD_THREADS=1 timeout 600 python tools/bench_c3.py --frames 400 --points-per-frame 500 --cpu-frames 0 > gpurun_out/c3_kd1.json 2>gpurun_out/c3_kd1.err
for t in 2 4 8; do SLAMGPU_KD_THREADS=$t timeout 600 python tools/bench_c3.py --frames 400 --points-per-frame 500 --cpu-frames 0 > gpurun_out/c3_kd$t.json 2>gpurun_out/c3_kd$t.err; done
python - <<'P'
import json
for t in (1,2,4,8):
    try:
        d=json.load(open(f'gpurun_out/c3_kd{t}.json')); print(t, round(d['frames_per_s']), d['ms_per_frame_parts'])
    except Exception as e: print(t,'failed',e)
P
nproc; lscpu | grep "Model name"
