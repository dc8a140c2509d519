#!/usr/bin/env python
"""Kernel time of the projection match kernel variants (walk / 8 / 32 lanes per point) on C2 and C4 shapes."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import racing_slam_b200 as pkg  # noqa: E402

syn = pkg.synthetic
cam = syn.KITTI
for name, walk, lanes in (("walk", "1", "0"), ("lanes32", "0", "32"), ("lanes8", "0", "8"), ("auto", "0", "0")):
    os.environ["SLAMGPU_PROJ_WALK"] = walk
    os.environ["SLAMGPU_PROJ_LANES"] = lanes
    with pkg.Context(0) as ctx:
        for label, npts, nkf in (("c2", 10000, 1), ("c4", 50000, 20), ("200k", 200000, 1)):
            sc = syn.projection_scene(npts, 2000, 4)
            rng = np.random.default_rng(4)
            ids, nkps, poses = [], [], []
            for k in range(nkf):
                ctx.frame_upload(k, sc.kp_xy, sc.kp_desc, None, sc.kp_matched)
                ids.append(k), nkps.append(2000)
                poses.append(sc.pose if k == 0 else syn.pose_colmajor(syn.make_pose(rng)))
            ctx.points_upload(0, sc.pos, sc.normal, sc.range, sc.obs_offset, sc.obs_desc)
            prm = pkg.ProjParams(cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["width"], cam["height"], 20.0, 64.0,
                                 0.5, 2.0, 1.25, 1 if nkf > 1 else 0)
            for _ in range(3):
                ctx.match_projection_batch(ids, nkps, 0, np.stack(poses), prm)
            ctx.profile_enable(True)
            ctx.profile_read(1)
            for _ in range(10):
                res = ctx.match_projection_batch(ids, nkps, 0, np.stack(poses), prm)
            ms, n = ctx.profile_read(1)
            ctx.profile_enable(False)
            print(f"{name:8s} {label:5s} kernel_us {ms / n * 1e3:9.1f}  matches {sum(int((r[0] >= 0).sum()) for r in res)}")
