#!/usr/bin/env python
"""Loop-closure query latency against a resident key-frame database (slamgpu_kfdb_*): the newest key frame against k
retrieved candidates of a 1000 x 4000 database, host wall time per call (pair list up, match lists down)."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import racing_slam_b200 as pkg  # noqa: E402

frames, rows = 1000, 4000
db = pkg.synthetic.keyframe_db_fast(frames, rows, 5)
rng = np.random.default_rng(1)
with pkg.Context(0) as ctx:
    for f in db:
        ctx.kfdb_append(3, f)
    out = (pkg.pinned_empty(64 * rows, np.int32), pkg.pinned_empty(64 * rows, np.int32))
    for k in (1, 5, 20, 64):
        cand = rng.choice(frames - 1, k, replace=False)
        pairs = np.stack([np.full(k, frames - 1), cand], 1).astype(np.int32)
        for _ in range(5):
            ctx.kfdb_match_pairs(3, pairs, out=out)
        t0 = time.perf_counter()
        reps = 100
        for _ in range(reps):
            o, ql, tl = ctx.kfdb_match_pairs(3, pairs, out=out)
        dt = (time.perf_counter() - t0) / reps
        print(f"query vs {k:2d} candidates of {frames} resident key frames: {dt * 1e6:8.1f} us per call, "
              f"{k * rows * rows / dt / 1e9:8.1f} G comparisons/s, {len(ql)} matches")
