#!/bin/bash
# MMA cost against the tile width N (timing experiment; variant 8 = MMAs without the epilogue, 12 = neither)
for n in variants N160 N128; do
  export SLAMGPU_DEBUG_LIBRARY=$PWD/racing-slam_b200/csrc/build/libslamgpu_debug_$n.so
  for v in 8 12; do
    echo -n "lib $n variant $v: "; SLAMGPU_TC4_TEAMS=1 SLAMGPU_TC_VARIANT=$v timeout 120 python tools/prof_tc.py 8192 4 | tail -1
  done
done
