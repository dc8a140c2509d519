#!/usr/bin/env python
"""Short driver for ncu: match_map on a 200k-point map into an 8000-keypoint 1080p frame (the C3 end state)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import racing_slam_b200 as pkg  # noqa: E402

syn = pkg.synthetic
cam = syn.HD1080
sc = syn.projection_scene(200000, 8000, 3, cam=cam)
with pkg.Context(0) as ctx:
    ctx.frame_upload(0, sc.kp_xy, sc.kp_desc, None, sc.kp_matched)
    ctx.points_upload(0, sc.pos, sc.normal, sc.range, sc.obs_offset, sc.obs_desc)
    prm = pkg.ProjParams(cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["width"], cam["height"], 20.0, 64.0, 0.5, 2.0,
                         1.25, 0)
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 6
    for _ in range(n):
        op, od = ctx.match_projection(0, 0, 8000, sc.pose, prm, sc.skip, None)
    print("matches", int((op >= 0).sum()))
