#!/bin/bash
# A/B of one environment switch on the SAME box (boxes differ by several per cent): tools/ab_bench.sh VAR val_a val_b [reps]
var=$1; a=$2; b=$3; reps=${4:-3}
for i in $(seq 1 $reps); do
  for v in $a $b; do
    env $var=$v timeout 300 python bench.py --no-also --steps 5 --warmup 3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline())
print('$var=$v', 'value %.0f' % d['value'], 'frac %.4f' % d['roofline']['frac'], 'launch_ms %.3f' % d['roofline']['avg_launch_ms'], 'e2e %.0f' % d['e2e']['value'], 'res %.0f' % d['e2e_resident_database']['value'], 'job %.4f' % d['c5_whole_job_strong']['seconds'], d['c5_whole_job_strong']['matches_total'], d['clocks']['sm_mhz'], d['clocks']['power_w_max'])
"
  done
done
