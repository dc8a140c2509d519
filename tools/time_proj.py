#!/usr/bin/env python
"""A/B timing of the projection match kernel (K4) variants on a B200: CUDA-event time per launch through the
library's profile hooks, plus a checksum of the results so that the variants can be compared for equality.

  python tools/time_proj.py            # runs every variant in a subprocess (env switches are read at context creation)
  python tools/time_proj.py --one NAME # one variant in this process (env already set)
"""
import hashlib
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

VARIANTS = {
    "v2_lanes8_cell32": {"SLAMGPU_PROJ_V2": "1", "SLAMGPU_PROJ_LANES": "8", "SLAMGPU_GRID_CELL": "32"},
    "v3_lanes8_cell32": {"SLAMGPU_PROJ_LANES": "8", "SLAMGPU_GRID_CELL": "32"},
    "v3_lanes8_cell16": {"SLAMGPU_PROJ_LANES": "8", "SLAMGPU_GRID_CELL": "16"},
    "v3_lanes8_cell12": {"SLAMGPU_PROJ_LANES": "8", "SLAMGPU_GRID_CELL": "12"},
    "v3_lanes32_cell16": {"SLAMGPU_PROJ_LANES": "32", "SLAMGPU_GRID_CELL": "16"},
    "fused_only": {"SLAMGPU_PROJ_TWOPASS": "2"},
    "fused_only_noqueue": {"SLAMGPU_PROJ_TWOPASS": "2", "SLAMGPU_PROJ_QUEUE": "0"},
    "two_pass_noqueue": {"SLAMGPU_PROJ_TWOPASS": "1", "SLAMGPU_PROJ_QUEUE": "0"},
    "two_pass_always": {"SLAMGPU_PROJ_TWOPASS": "1"},
    "default": {},
}


def one(name):
    import racing_slam_b200 as pkg
    syn = pkg.synthetic
    kitti, hd = syn.KITTI, syn.HD1080

    def prm(cam, replace):
        return pkg.ProjParams(cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["width"], cam["height"], 20.0, 64.0, 0.5,
                              2.0, 1.25, replace)

    with pkg.Context(0) as ctx:
        def timed(fn, reps=30):
            for _ in range(3):
                r = fn()
            ctx.profile_enable(True)
            ctx.profile_read(1)
            for _ in range(reps):
                r = fn()
            ms, n = ctx.profile_read(1)
            ctx.profile_enable(False)
            return ms / max(n, 1) * 1e3, r

        def digest(arrs):
            h = hashlib.sha1()
            for a in arrs:
                h.update(np.ascontiguousarray(a).tobytes())
            return h.hexdigest()[:10]

        sc = syn.projection_scene(10000, 2000, 2)
        ctx.frame_upload(0, sc.kp_xy, sc.kp_desc, None, sc.kp_matched)
        ctx.points_upload(0, sc.pos, sc.normal, sc.range, sc.obs_offset, sc.obs_desc)
        us, r = timed(lambda: ctx.match_projection(0, 0, 2000, sc.pose, prm(kitti, 0), sc.skip, None))
        print(f"{name:20s} c2    kernel_us {us:9.1f}  matches {int((r[0] >= 0).sum())}  sha {digest(r)}")

        sc4 = syn.projection_scene(50000, 2000, 4)
        rng = np.random.default_rng(4)
        ids, poses = [], []
        for k in range(20):
            xy = syn.keypoints_integer(2000, kitti["width"], kitti["height"], 700 + k)
            ctx.frame_upload(100 + k, xy, sc4.kp_desc[rng.permutation(2000)], None, None)
            ids.append(100 + k)
            poses.append(syn.pose_colmajor(syn.make_pose(rng)))
        ctx.points_upload(1, sc4.pos, sc4.normal, sc4.range, sc4.obs_offset, sc4.obs_desc)
        poses = np.stack(poses)
        us, r = timed(lambda: ctx.match_projection_batch(ids, [2000] * 20, 1, poses, prm(kitti, 1)), 10)
        print(f"{name:20s} c4    kernel_us {us:9.1f}  matches {sum(int((a >= 0).sum()) for a, _ in r)}  "
              f"sha {digest([x for pr in r for x in pr])}")

        sc3 = syn.projection_scene(200000, 8000, 3, cam=hd)
        ctx.frame_upload(2, sc3.kp_xy, sc3.kp_desc, None, sc3.kp_matched)
        ctx.points_upload(2, sc3.pos, sc3.normal, sc3.range, sc3.obs_offset, sc3.obs_desc)
        us, r = timed(lambda: ctx.match_projection(2, 2, 8000, sc3.pose, prm(hd, 0), sc3.skip, None), 10)
        print(f"{name:20s} 200k  kernel_us {us:9.1f}  matches {int((r[0] >= 0).sum())}  sha {digest(r)}")


if __name__ == "__main__":
    if len(sys.argv) > 2 and sys.argv[1] == "--one":
        one(sys.argv[2])
    else:
        print("# tools/time_proj.py on B200: proj match kernel, CUDA-event time per launch; c2 = 10k points x 1 frame "
              "(KITTI, 2000 kp), c4 = 50k points x 20 key frames, 200k = 200k points x 1 frame (1080p, 8000 kp)")
        for name, env in VARIANTS.items():
            e = dict(os.environ)
            for k in ("SLAMGPU_PROJ_V2", "SLAMGPU_PROJ_LANES", "SLAMGPU_GRID_CELL", "SLAMGPU_PROJ_WALK", "SLAMGPU_PROJ_TWOPASS", "SLAMGPU_PROJ_QUEUE"):
                e.pop(k, None)
            e.update(env)
            sys.stdout.flush()
            subprocess.run([sys.executable, os.path.abspath(__file__), "--one", name], env=e, check=False)
