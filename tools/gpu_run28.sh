#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_bf.py -m gpu -x -q > gpurun_out/r28_tests.log 2>&1; echo "tests rc=$?"; tail -2 gpurun_out/r28_tests.log
for v in 256 0 256 0; do
SLAMGPU_TC_VARIANT=$v timeout 300 python bench.py --no-also --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('variant $v', round(d['value'],1), round(d['ms_per_step'],3), round(d['roofline']['frac'],4), round(d['e2e']['value'],1))"
done
