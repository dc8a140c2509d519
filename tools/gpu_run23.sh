#!/bin/bash
# two-pass K4: parity, then A/B timing
timeout 900 python -m pytest tests/test_gpu_proj.py tests/test_gpu_track.py tests/test_adapter.py -m gpu -x -q > gpurun_out/r23_tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/r23_tests.log
timeout 600 python tools/time_proj.py > gpurun_out/r23_time_proj.txt 2>&1
grep -v "^#" gpurun_out/r23_time_proj.txt | grep "fused\|two_pass\|default"
