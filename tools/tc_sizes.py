#!/usr/bin/env python
"""Device-resident call time of match_descriptors / crosscheck by size for both brute-force engines."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import racing_slam_b200 as pkg  # noqa: E402

ctxs = {}
for eng in ("int", "tc"):
    os.environ["SLAMGPU_BF_ENGINE"] = eng
    ctxs[eng] = pkg.Context(0)
for n in (256, 512, 1000, 2000, 3000, 4000, 6000, 8000, 16000):
    q, t = pkg.synthetic.bf_correlated(n, n, n)
    dq, dt = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
    out = torch.empty(n, dtype=torch.int32, device="cuda")
    out2 = torch.empty(n, dtype=torch.int32, device="cuda")
    cnt = torch.empty(1, dtype=torch.int32, device="cuda")
    line = f"n {n:6d}"
    for eng, ctx in ctxs.items():
        st = torch.cuda.ExternalStream(ctx.stream)
        for what, fn in (("match_desc", lambda: ctx.match_descriptors_dev(dq.data_ptr(), n, dt.data_ptr(), n, 64.0, 0.75,
                                                                         out.data_ptr(), cnt.data_ptr())),
                         ("crosscheck", lambda: ctx.crosscheck_dev(dq.data_ptr(), n, dt.data_ptr(), n, out.data_ptr(),
                                                                   out2.data_ptr()))):
            for _ in range(3):
                fn()
            ctx.sync()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(20):
                fn()
            e1.record(st)
            torch.cuda.synchronize()
            line += f"  {eng}/{what} {e0.elapsed_time(e1) * 50.0:8.1f} us"
    print(line, flush=True)
