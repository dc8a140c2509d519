#!/bin/bash
cd $GRAFT_REPO_ROOT
for i in 1 2; do
for v in 1 0; do
  SLAMGPU_PROJ_PREFETCH=$v timeout 300 python tools/time_proj.py --one prefetch$v 2>&1 | grep kernel_us
done
done
