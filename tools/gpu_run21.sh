cd $GRAFT_REPO_ROOT
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest_gpu.log
tail -n 3 gpurun_out/pytest_gpu.log
SLAMGPU_KD_ASYNC=0 timeout 600 python tools/bench_c3.py --frames 400 --points-per-frame 500 --cpu-frames 0 > gpurun_out/c3_sync.json 2>gpurun_out/c3_sync.err
timeout 600 python tools/bench_c3.py --frames 400 --points-per-frame 500 --cpu-frames 0 > gpurun_out/c3_async.json 2>gpurun_out/c3_async.err
timeout 900 python tools/bench_c3.py --frames 2000 > gpurun_out/c3_full.json 2> gpurun_out/c3_full.err
python - <<'P'
import json
for t in ('sync','async','full'):
    try:
        d=json.load(open(f'gpurun_out/c3_{t}.json')); print(t, round(d['frames_per_s']), d['ms_per_frame_parts'], d.get('cpu',{}).get('gpu_results_identical_on_sample'))
    except Exception as e: print(t,'failed',e)
P
