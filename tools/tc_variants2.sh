#!/bin/bash
# timing-experiment build: parts of the FP4 scan switched off (4 = no MMAs, 8 = no TMEM reads / fold, 128 = no scan arithmetic)
export SLAMGPU_DEBUG_LIBRARY=$PWD/racing-slam_b200/csrc/build/libslamgpu_debug_variants.so
for t in ${TEAMS_LIST:-0 1}; do
for v in ${VARIANTS:-0 132 128 4 8 12}; do
    echo -n "teams $t variant $v: "; SLAMGPU_TC4_TEAMS=$t SLAMGPU_TC_VARIANT=$v timeout 120 python tools/prof_tc.py 8192 4 | tail -1
done; done
