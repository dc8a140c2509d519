#!/bin/bash
# C3 full sequence + ncu of the two K4 passes (after the queue / binary-search changes)
cd $GRAFT_REPO_ROOT
timeout 600 python tools/bench_c3.py --frames 2000 --cpu-frames 6 > gpurun_out/c3_v4.json 2> gpurun_out/c3_v4.err; echo "c3 rc=$?"
python - <<'P'
import json
d=json.loads(open('gpurun_out/c3_v4.json').read().strip().splitlines()[-1])
print(d['frames_per_s'], d['ms_per_frame_parts'], d['ms_worst_frame_parts'], d['proj_match_kernel_us_avg'])
print(d.get('cpu'))
P
timeout 600 ncu --set full --clock-control none --import-source on -k regex:proj_gate_kernel\|proj_search_kernel -s 6 -c 2 \
  -o gpurun_out/proj_twopass_v5 -f python tools/prof_proj_big.py 6 > gpurun_out/ncu_twopass_v5.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/proj_twopass_v5.ncu-rep --page details > gpurun_out/proj_twopass_v5_details.txt 2>&1
grep -n "proj_gate_kernel\|proj_search_kernel\|  Duration\|Issue Slots Busy\|Executed Ipc Active\|Registers Per\|Achieved Occupancy\|Executed Instructions  \|Avg. Active Threads" gpurun_out/proj_twopass_v5_details.txt | head -30
