#!/bin/bash
# full validation after the two-pass K4 + ncu of its two kernels on the 200k map
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r24_tests.log 2>&1; echo "tests rc=$?"
tail -2 gpurun_out/r24_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r24_smoke.log 2>&1; echo "smoke rc=$?"
timeout 900 python bench.py > gpurun_out/r24_bench.json 2> gpurun_out/r24_bench.err; echo "bench rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:proj_gate_kernel\|proj_search_kernel -s 6 -c 2 \
  -o gpurun_out/proj_twopass -f python tools/prof_proj_big.py 6 > gpurun_out/ncu_twopass.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/proj_twopass.ncu-rep --page details > gpurun_out/proj_twopass_details.txt 2>&1
grep -n "proj_gate_kernel\|proj_search_kernel\|Duration\|Issue Slots Busy\|Executed Ipc Active\|Registers Per\|Achieved Occupancy\|DRAM Throughput" gpurun_out/proj_twopass_details.txt | head -30
