#!/usr/bin/env python
"""Short driver for ncu / timing: RANSAC inlier scoring of 1000 problems x 200 models, device-resident."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import racing_slam_b200 as pkg  # noqa: E402
import torch  # noqa: E402

syn = pkg.synthetic
n_prob, n_mod = 1000, 200
rng = np.random.default_rng(9)
sizes = [int(x) for x in rng.integers(40, 560, n_prob)]
objs, imgs, models = syn.pnp_problems(2, sizes, [n_mod] * n_prob)
obj, img, rt = np.concatenate(objs), np.concatenate(imgs), np.concatenate(models)
cnt = np.array(sizes, np.int32)
begin = (np.cumsum(cnt) - cnt).astype(np.int32)
hp = np.repeat(np.arange(n_prob, dtype=np.int32), n_mod)
evals = int(cnt.astype(np.int64).sum()) * n_mod
with pkg.Context(0) as ctx:
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    d_obj, d_img, d_rt, d_begin, d_cnt, d_hp = d(obj), d(img), d(rt), d(begin), d(cnt), d(hp)
    d_good = torch.empty(len(hp), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
    ctx.profile_enable(True)
    for _ in range(reps):
        ctx.pnp_score_dev(d_obj.data_ptr(), d_img.data_ptr(), d_begin.data_ptr(), d_cnt.data_ptr(), n_prob,
                          d_rt.data_ptr(), d_hp.data_ptr(), len(hp), syn.PNP_K4, 4.0, d_good.data_ptr())
    ms, n = ctx.profile_read(3)
    print(f"pnp_score_kernel: {ms / n * 1e3:.1f} us per launch, {evals / (ms / n * 1e-3) / 1e9:.1f} G evaluations/s, "
          f"{evals} evaluations, inliers total {int(d_good.sum())}")
