#!/bin/bash
# Where does the FP4 scan's time go?  The timing-experiment build (make -C racing-slam_b200/csrc debug) with parts of
# the kernel switched off: 4 = no MMAs, 8 = no TMEM reads / fold, 128 = no scan arithmetic, 32 = no rare path.
export SLAMGPU_DEBUG_LIBRARY=$PWD/racing-slam_b200/csrc/build/libslamgpu_debug_variants.so
for v in 0 4 8 12 128 32 160; do
  for n in 0 1; do
    echo -n "variant $v narrow $n: "; SLAMGPU_TC4_NARROW=$n SLAMGPU_TC_VARIANT=$v timeout 120 python tools/prof_tc.py 8192 4 | tail -1
  done
done
