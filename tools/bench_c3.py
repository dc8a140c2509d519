"""BASELINE configs[2]: 1080p tracking sequence, 8000 ORB features per frame, frame-to-frame
cross-check matching + the tracking pair against a map growing 0 -> 200k points, 1 GPU, through the C ABI
with HOST buffers (every copy inside the timed region).

Per frame, in the order the reference's main loop does it (src/Tracker.cpp:66-88, src/Mapper.cpp:152-174):
  1. Frame construction: upload; the KD-tree builds on a host thread    slamgpu_frame_upload
  2. match_features(prev, cur)            8000 x 8000 cross-check      slamgpu_match_features
  3. match_with_last_key_frame + match_with_map                         slamgpu_track_match
  4. Mapper::insert adds this frame's new points                        slamgpu_points_append
Frame synthesis is outside the timer.  The CPU arm (bench.py: c3_cpu_arm) runs the same steps on a bounded
sample of frames (cv2 BFMatcher cross-check when cv2 is present, else the oracle; the oracle's KD build and two
projection searches) and every sampled frame doubles as a parity check of the GPU results.

  python tools/bench_c3.py [--frames 2000] [--points-per-frame 100] [--cpu-frames 6]
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import gc
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

MAX_DIST, RADIUS = 64.0, 20.0


def _params(pkg, cam):
    return pkg.ProjParams(cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["width"], cam["height"], RADIUS, MAX_DIST,
                          0.5, 2.0, 1.25, 0)


def run_sequence(ctx, pkg, n_frames=2000, nkp=8000, ppf=100, seed=3, cpu_frames=6, cpu_arm=None, frame_slot=700,
                 set_slot=700):
    """-> dict for the bench line.  cpu_arm(pkg, seq, kept, prm) is bench.py's host leg (the only place the
    oracle is executed).  Timed with the host clock around the four C-ABI calls of each frame (they
    return after their device work and copies have completed)."""
    syn = pkg.synthetic
    seq = syn.TrackingSequence(n_frames, nkp, ppf, seed)
    prm = _params(pkg, seq.cam)
    empty = seq.pool
    ctx.points_upload(set_slot, empty.pos[:0], empty.normal[:0], empty.range[:0], np.zeros(1, np.int32),
                      empty.obs_desc[:0])
    sample = set()
    if cpu_frames and cpu_arm is not None:
        sample = set(np.linspace(n_frames // 10, n_frames - 1, cpu_frames).astype(int).tolist())
    t_parts = np.zeros(4)
    t_worst = np.zeros(4)
    t_all = []
    gc_was_on = gc.isenabled()
    gc.disable()              # the harness's own collector pauses (milliseconds) are not the library's time
    kept = {}
    prev = seq.frame(0)
    n_kf = n_map = 0
    ctx.profile_enable(True)
    ctx.profile_read(1)
    for f in range(n_frames):
        cur = seq.frame(f)
        new = seq.new_points(f)
        t0 = time.perf_counter()
        # Frame construction first, as in the reference (the Frame exists before anything is matched): the
        # upload returns once the descriptors are resident and the KD-tree build continues on a host thread
        # of the library while the cross-check match runs on the GPU; track_match waits for it.
        ctx.frame_upload(frame_slot, cur["kp_xy"], cur["kp_desc"], None, cur["kp_matched"])
        t1 = time.perf_counter()
        ft, fq = ctx.match_features(prev["kp_desc"], cur["kp_desc"], MAX_DIST)
        t2 = time.perf_counter()
        res = ctx.track_match(frame_slot, set_slot, nkp, cur["pose"], prm, cur["matched_points"], cur["kf_points"])
        t3 = time.perf_counter()
        ctx.points_append(set_slot, *new)
        t4 = time.perf_counter()
        t_parts += (t1 - t0, t2 - t1, t3 - t2, t4 - t3)
        t_worst = np.maximum(t_worst, (t1 - t0, t2 - t1, t3 - t2, t4 - t3))
        t_all.append((t1 - t0, t2 - t1, t3 - t2, t4 - t3))
        n_kf += int((res[0] >= 0).sum())
        n_map += int((res[2] >= 0).sum())
        if f in sample:
            kept[f] = (prev, cur, (ft, fq), res)
        prev = cur
    if gc_was_on:
        gc.enable()
    k_ms, k_n = ctx.profile_read(1)
    ctx.profile_enable(False)
    spec_calls, spec_repeats = ctx.speculation_count()
    total = float(t_parts.sum())
    out = {
        "workload": f"{n_frames} frames x {nkp} keypoints (1920x1080), map 0 -> {n_frames * ppf} points "
                    f"({ppf}/frame), radius {RADIUS:g}, per frame: frame_upload (KD build on a host thread) + match_features + "
                    "track_match (key-frame subset then whole map) + points_append",
        "frames_per_s": n_frames / total, "ms_per_frame": total / n_frames * 1e3,
        "ms_per_frame_parts": {"frame_upload": t_parts[0] / n_frames * 1e3, "match_features": t_parts[1] / n_frames * 1e3,
                               "track_match_incl_kd_wait": t_parts[2] / n_frames * 1e3,
                               "points_append": t_parts[3] / n_frames * 1e3},
        # the headline is the mean (whole run / frames); the medians show what a frame costs when the host is not
        # disturbed (a single hiccup of the shared host -- tens of ms -- moves the mean of a short run)
        "ms_median_frame_parts": dict(zip(("frame_upload", "match_features", "track_match_incl_kd_wait", "points_append"),
                                          (float(x) * 1e3 for x in np.median(np.asarray(t_all), axis=0)))),
        "ms_median_frame": float(np.median(np.asarray(t_all).sum(axis=1))) * 1e3,
        "ms_worst_frame_parts": {"frame_upload": t_worst[0] * 1e3, "match_features": t_worst[1] * 1e3,
                                 "track_match_incl_kd_wait": t_worst[2] * 1e3, "points_append": t_worst[3] * 1e3},
        "gcmp_s_bruteforce_part": nkp * nkp / (t_parts[1] / n_frames) / 1e9,
        "proj_match_kernel_us_avg": k_ms / max(k_n, 1) * 1e3, "proj_match_kernel_launches": int(k_n),
        "keyframe_matches": n_kf, "map_matches": n_map, "copies": "inside", "clock": "host (Python gc off during the timed loop)",
        "speculative_calls_so_far": spec_calls, "speculative_calls_repeated_with_the_tree": spec_repeats,
    }
    if kept:
        out["cpu"] = cpu_arm(pkg, seq, kept, prm)
    ctx.frame_release(frame_slot)
    ctx.points_release(set_slot)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=2000)
    ap.add_argument("--points-per-frame", type=int, default=100)
    ap.add_argument("--keypoints", type=int, default=8000)
    ap.add_argument("--cpu-frames", type=int, default=6)
    a = ap.parse_args()
    import racing_slam_b200 as pkg
    with pkg.Context(0) as ctx:
        import bench
        run_sequence(ctx, pkg, 20, a.keypoints, a.points_per_frame, 3, 0)          # warm-up
        out = run_sequence(ctx, pkg, a.frames, a.keypoints, a.points_per_frame, 3, a.cpu_frames, bench.c3_cpu_arm)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
