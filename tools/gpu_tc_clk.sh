cd $GRAFT_REPO_ROOT
for v in 0 4 8 2 16; do
  nvidia-smi --query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap,clocks_event_reasons.hw_slowdown,clocks_event_reasons.sw_thermal_slowdown --format=csv,noheader,nounits -lms 100 > gpurun_out/clk_v$v.csv &
  SMI=$!
  echo "variant $v"; SLAMGPU_TC_VARIANT=$v timeout 200 python tools/prof_tc.py 2048 300
  kill $SMI
  echo "clock samples (sorted by power, top 12):"; sort -t, -k2 -n -r gpurun_out/clk_v$v.csv | head -12 | tr '\n' ';'; echo
done > gpurun_out/tc_clk.log 2>&1
cat gpurun_out/tc_clk.log
