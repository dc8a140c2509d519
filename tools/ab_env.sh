#!/bin/bash
# Same-box A/B of one environment switch on the single-call latencies (tools/call_latency), alternating, three rounds:
#   tools/ab_env.sh SLAMGPU_BF_INLINE_COPY=0
python tools/run_call_latency.py --keep > /dev/null 2>&1
for r in 1 2 3; do
  for v in default "$1"; do
    echo "== $v round $r"
    if [ "$v" = default ]; then e=""; else e="$v"; fi
    env $e tools/call_latency gpurun_out/call_latency_inputs.blob 300 | python -c "
import json,sys
d=json.load(sys.stdin)
print('  '.join(f\"{k.replace('_candidates','').replace('match_','').replace('projection','proj')}={v['us_per_call']:.1f}\" for k,v in d.items() if isinstance(v,dict)))"
  done
done
