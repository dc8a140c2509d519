#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_bf.py tests/test_gpu_proj.py tests/test_gpu_track.py -m gpu -x -q > gpurun_out/r25_tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/r25_tests.log
timeout 900 python bench.py > gpurun_out/r25_bench.json 2> gpurun_out/r25_bench.err; echo "bench rc=$?"
python - <<'P'
import json
d=json.loads(open('gpurun_out/r25_bench.json').read().strip().splitlines()[-1])
print({k:d[k] for k in ('value','ms_per_step','e2e','gpu_launches','clocks')})
c=d['also']['c3_tracking_sequence_200_frames']; print(c['frames_per_s'], c['ms_per_frame_parts'], c.get('ms_worst_frame_parts'))
print(d['also']['c4_fuse_batch_20_keyframes_x_50k_points'])
P
