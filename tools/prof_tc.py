#!/usr/bin/env python
"""Short driver for ncu: the C5-shaped batch (N pairs of 4000 x 4000) through the tensor-core scan."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("SLAMGPU_BF_ENGINE", "tc")
import torch  # noqa: E402

import racing_slam_b200 as pkg  # noqa: E402

npairs = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
nf, nd = 1000, 4000
db = pkg.synthetic.keyframe_db_fast(nf, nd, 5)
off = (np.arange(nf + 1, dtype=np.int64) * nd).astype(np.int32)
pairs = np.ascontiguousarray(pkg.sharding.all_pairs(nf)[:npairs].astype(np.int32))
with pkg.Context(0) as ctx:
    d_db = torch.from_numpy(db.reshape(-1, 32)).cuda()
    d_off = torch.from_numpy(off).cuda()
    d_pairs = torch.from_numpy(pairs).cuda()
    d_out = torch.empty((npairs, nd), dtype=torch.int32, device="cuda")
    d_cnt = torch.empty(npairs, dtype=torch.int32, device="cuda")
    st = torch.cuda.ExternalStream(ctx.stream)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for r in range(reps):
        if r == reps - 1:
            e0.record(st)
        ctx.match_descriptors_batch_dev(d_db.data_ptr(), d_off.data_ptr(), nf, d_pairs.data_ptr(), npairs, nd, 64.0, 0.75,
                                        d_out.data_ptr(), d_cnt.data_ptr())
    e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(f"pairs {npairs}  last call {ms:.3f} ms  {npairs * 16e6 / ms / 1e6:.1f} Gcmp/s  matches {int(d_cnt.sum())}")
