#!/bin/bash
# Same-box A/B of the single-call latencies: tools/call_latency against build_ab/base/libslamgpu.so (a build of an
# earlier commit) and against the tree's library, alternating, three rounds.  Run on the GPU box:
#   tools/ab_calls.sh > gpurun_out/ab_calls.txt
python tools/run_call_latency.py --keep > /dev/null 2>&1
for r in 1 2 3; do
  for v in base new; do
    if [ $v = base ]; then export LD_LIBRARY_PATH=$PWD/build_ab/base; else unset LD_LIBRARY_PATH; fi
    echo "== $v round $r $EXTRA_ENV"
    env $EXTRA_ENV tools/call_latency gpurun_out/call_latency_inputs.blob 300 | python -c "
import json,sys
d=json.load(sys.stdin)
print('  '.join(f\"{k.replace('_candidates','').replace('match_','').replace('projection','proj')}={v['us_per_call']:.1f}\" for k,v in d.items() if isinstance(v,dict)))"
  done
done
