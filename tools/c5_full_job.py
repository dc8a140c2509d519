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
# under torchrun: the pair list sharded block-cyclically over the ranks (strong scaling, no data-path collective;
# gloo carries the barrier, the max of the times and the sum of the match counts)
world, rank, local = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
if world > 1:
    import torch
    import torch.distributed as dist
    dist.init_process_group("gloo")
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
n_all = len(pairs)
pairs = pkg.sharding.shard_pairs(pairs, rank, world, block=64)
resident = len(sys.argv) > 3 and sys.argv[3] == "resident"       # key frames appended once (slamgpu_kfdb_*)
with pkg.Context(local) as ctx:
    ctx.match_descriptors_batch(db, off, pairs[:2048], 64.0, 0.75)          # warm-up: allocations, expanded copy
    t_up = time.perf_counter()
    if resident:
        for f in range(frames):
            ctx.kfdb_append(7, db[f * rows:(f + 1) * rows])
        ctx.kfdb_match_pairs(7, pairs[:2048], out=out, capacity=2048 * rows)
    if world > 1:
        dist.barrier()
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
if world > 1:
    t = torch.tensor([dt], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    m = torch.tensor([total, per_pair_max], dtype=torch.int64)
    mx = m.clone()
    dist.all_reduce(m, op=dist.ReduceOp.SUM)
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    dt, total, per_pair_max = float(t[0]), int(m[0]), int(mx[1])
    dist.destroy_process_group()
    if rank != 0:
        sys.exit(0)
cmp_total = float(n_all) * rows * rows
print(json.dumps({"workload": "C5 whole job: 499,500 ordered pairs of 4000x4000, host buffers, copies inside",
                  "pairs": n_all, "n_gpus": world, "slice_pairs": step, "seconds": dt, "gcmp_s": cmp_total / dt / 1e9,
                  "matches_total": total, "matches_max_per_pair": per_pair_max, "host_memory": "pinned" if pin else "pageable", "database": "resident (appended once, %.3f s incl. one warm-up call)" % t_up if resident else "uploaded per call", "clock": "host wall"}))
