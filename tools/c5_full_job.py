#!/usr/bin/env python
"""BASELINE configs[4] as one whole job: every ordered pair (query > candidate) of 1000 key frames x 4000 descriptors
= 499,500 pairs = 7.99e12 Hamming comparisons, through the host-buffer batch call (slamgpu_match_descriptors_batch)
in slices of 32768 pairs.  Prints one JSON line: wall seconds, comparisons/s, total matches (a checksum that must not
depend on the slice size or on how the pairs are sharded)."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import racing_slam_b200 as pkg  # noqa: E402

frames, rows = 1000, 4000
step = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
pin = len(sys.argv) <= 2 or sys.argv[2] != "pageable"
db = pkg.synthetic.keyframe_db_fast(frames, rows, 5).reshape(-1, 32)
if pin:                                   # page-locked descriptor set and result arrays (slamgpu_host_alloc)
    h = pkg.pinned_empty(db.shape, np.uint8)
    h[...] = db
    db = h
    out = (pkg.pinned_empty(step * 1024, np.int32), pkg.pinned_empty(step * 1024, np.int32))
else:
    out = None
off = (np.arange(frames + 1, dtype=np.int64) * rows).astype(np.int32)
pairs = pkg.sharding.all_pairs(frames)
resident = len(sys.argv) > 3 and sys.argv[3] == "resident"       # key frames appended once (slamgpu_kfdb_*)
with pkg.Context(0) as ctx:
    ctx.match_descriptors_batch(db, off, pairs[:2048], 64.0, 0.75)          # warm-up: allocations, expanded copy
    t_up = time.perf_counter()
    if resident:
        for f in range(frames):
            ctx.kfdb_append(7, db[f * rows:(f + 1) * rows])
        ctx.kfdb_match_pairs(7, pairs[:2048], out=out, capacity=2048 * rows)
    t0 = time.perf_counter()
    t_up = t0 - t_up
    total, per_pair_max = 0, 0
    for s in range(0, len(pairs), step):
        if resident:
            o, ql, tl = ctx.kfdb_match_pairs(7, pairs[s:s + step], 64.0, 0.75, out=out, capacity=step * 1024)
        else:
            o, ql, tl = ctx.match_descriptors_batch(db, off, pairs[s:s + step], 64.0, 0.75, out=out)
        total += len(ql)
        per_pair_max = max(per_pair_max, int(np.diff(o).max()))
    dt = time.perf_counter() - t0
cmp_total = float(len(pairs)) * rows * rows
print(json.dumps({"workload": "C5 whole job: 499,500 ordered pairs of 4000x4000, host buffers, copies inside",
                  "pairs": len(pairs), "slice_pairs": step, "seconds": dt, "gcmp_s": cmp_total / dt / 1e9,
                  "matches_total": total, "matches_max_per_pair": per_pair_max, "host_memory": "pinned" if pin else "pageable", "database": "resident (appended once, %.3f s incl. one warm-up call)" % t_up if resident else "uploaded per call", "clock": "host wall"}))
