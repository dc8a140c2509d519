#!/bin/bash
# like ab_bench.sh, but keeps every run's stderr and prints the failure instead of a JSON decode error
var=$1; a=$2; b=$3; reps=${4:-3}
for i in $(seq 1 $reps); do
  for v in $a $b; do
    env $var=$v timeout 300 python bench.py --no-also --steps 5 --warmup 3 > gpurun_out/ab_run.json 2> gpurun_out/ab_run.err
    rc=$?
    if [ $rc -ne 0 ]; then echo "$var=$v FAILED rc=$rc: $(tail -1 gpurun_out/ab_run.err | cut -c1-200)"; continue; fi
    python - "$var=$v" <<'PY'
import json,sys
d=json.loads(open('gpurun_out/ab_run.json').readline())
print(sys.argv[1], 'value %.0f' % d['value'], 'frac %.4f' % d['roofline']['frac'], 'launch_ms %.3f' % d['roofline']['avg_launch_ms'], 'e2e %.0f' % d['e2e']['value'], 'res %.0f' % d['e2e_resident_database']['value'], 'job %.4f' % d['c5_whole_job_strong']['seconds'], d['c5_whole_job_strong']['matches_total'], d['parity_sample']['identical'], d['clocks']['sm_mhz'], d['clocks']['power_w_max'])
PY
  done
done
