#!/bin/bash
# One parametrised GPU runner (replaces the per-lease gpu_run*.sh scripts of round 1):
#   tools/gpu.sh <timeout-seconds> '<command run on the B200 box>'
# retries while the pod answers "busy" (exit 3), logs to gpurun_out/local_<tag>.log
t=${1:-900}; shift
for i in $(seq 1 40); do
  /usr/local/graft/bin/gpurun --timeout "$t" -- "$@"
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 90
done
exit 3
