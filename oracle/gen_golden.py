#!/usr/bin/env python
"""Generate tests/golden/*.npz -- run in the BUILD container only (needs cv2 and /root/reference).

TEST INFRASTRUCTURE ONLY.  What gets pinned:

  bf_*.npz   inputs + outputs of cv2.BFMatcher (OpenCV 4.13.0), the same library code path the
             reference calls at src/MapMatcher.cpp:147-148 (knnMatch, k=2) and
             src/features/FeatureExtractor.cpp:20-21 (crossCheck=True, k=1), plus the
             reference's own filters applied to those DMatch rows (src/MapMatcher.cpp:151-161,
             src/features/FeatureExtractor.cpp:24-27).
  kd_*.npz   inputs + outputs of the REFERENCE's KDTree2D::radius_search
             (/root/reference/src/KDTree.cpp compiled by oracle/Makefile into oracle/_ref/).
  proj_*.npz inputs + outputs of the C++ oracle's projection-guided search (restatement;
             float part parity-unpinned, see oracle/slam_oracle.cpp header).  Regression pins.

Usage:  python oracle/gen_golden.py            (writes tests/golden/, prints a summary)
"""
from __future__ import annotations

import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import bf_model, cpu_oracle  # noqa: E402

spec = importlib.util.spec_from_file_location("synthetic", os.path.join(ROOT, "racing-slam_b200", "synthetic.py"))
synthetic = importlib.util.module_from_spec(spec)
sys.modules["synthetic"] = synthetic
spec.loader.exec_module(synthetic)

GOLD = os.path.join(ROOT, "tests", "golden")


def cv2_knn2(q, t):
    import cv2
    idx = np.full((len(q), 2), -1, np.int32)
    dist = np.full((len(q), 2), -1, np.int32)
    if len(q) == 0 or len(t) == 0:
        return idx, dist
    k = 2 if len(t) >= 2 else 1                                  # src/MapMatcher.cpp:145
    rows = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q, t, k=k)
    for i, row in enumerate(rows):
        for j, m in enumerate(row):
            assert m.queryIdx == i and m.imgIdx == 0
            assert float(m.distance) == int(m.distance)
            idx[i, j] = m.trainIdx
            dist[i, j] = int(m.distance)
    return idx, dist


def cv2_crosscheck(q, t):
    import cv2
    idx = np.full(len(q), -1, np.int32)
    dist = np.full(len(q), -1, np.int32)
    if len(q) == 0 or len(t) == 0:
        return idx, dist
    rows = cv2.BFMatcher(cv2.NORM_HAMMING, True).knnMatch(q, t, k=1)
    assert len(rows) == len(q)
    for i, row in enumerate(rows):
        if len(row):
            assert row[0].queryIdx == i
            idx[i] = row[0].trainIdx
            dist[i] = int(row[0].distance)
    return idx, dist


def ref_filter_match_descriptors(idx, dist, nt, max_dist, ratio):
    """src/MapMatcher.cpp:151-161 applied to cv2's rows."""
    oq, ot = [], []
    k = 2 if nt >= 2 else 1
    for i in range(len(idx)):
        if idx[i, 0] < 0 or np.float32(dist[i, 0]) > np.float32(max_dist):
            continue
        if k > 1 and idx[i, 1] >= 0 and np.float32(dist[i, 0]) > np.float32(ratio) * np.float32(dist[i, 1]):
            continue
        oq.append(i)
        ot.append(idx[i, 0])
    return np.array(oq, np.int32), np.array(ot, np.int32)


def ref_filter_match_features(idx, dist, max_dist):
    """src/features/FeatureExtractor.cpp:24-27 applied to cv2's cross-check rows."""
    keep = (idx >= 0) & (dist.astype(np.float32) < np.float32(max_dist))
    qi = np.nonzero(keep)[0].astype(np.int32)
    return idx[qi], qi


def bf_cases():
    yield "uniform_64x48", synthetic.bf_uniform(64, 48, 1)
    yield "correlated_200x150", synthetic.bf_correlated(200, 150, 2)
    yield "ties_96x80", synthetic.bf_tie_heavy(96, 80, 3)
    yield "ties_300x257", synthetic.bf_tie_heavy(300, 257, 4, pool=12)
    yield "nt1_10x1", synthetic.bf_correlated(10, 1, 5)
    yield "nt2_10x2", synthetic.bf_correlated(10, 2, 6)
    yield "nq1_1x33", synthetic.bf_correlated(1, 33, 7)
    yield "correlated_512x384", synthetic.bf_correlated(512, 384, 8, max_flip=90)
    q, t = synthetic.bf_correlated(130, 130, 9, max_flip=0, frac_replaced=0.0)   # exact copies: d0 == 0
    yield "exactcopies_130x130", (q, t)
    # threshold edges: distances of exactly 64 (passes match_descriptors '>' gate, fails the
    # match_features '<' gate) and exact 0.75 ratios (48 vs 64)
    rng = np.random.default_rng(10)
    q = rng.integers(0, 256, (40, 32), dtype=np.uint8)
    t = np.concatenate([synthetic.flip_bits(q, 64, rng), synthetic.flip_bits(q, 48, rng),
                        synthetic.flip_bits(q, 63, rng)[:10]])
    yield "edges_40x90", (q, t)


def main():
    os.makedirs(GOLD, exist_ok=True)
    cpu_oracle.build()
    n_bf = 0
    for name, (q, t) in bf_cases():
        k_idx, k_dist = cv2_knn2(q, t)
        c_idx, c_dist = cv2_crosscheck(q, t)
        # the model and the C++ oracle must agree with cv2 before anything is written
        for impl in (bf_model, cpu_oracle):
            mi, md = impl.knn2(q, t)
            assert np.array_equal(mi, k_idx) and np.array_equal(md, k_dist), (name, impl.__name__, "knn2")
            ci, cd = impl.crosscheck(q, t)
            assert np.array_equal(ci, c_idx) and np.array_equal(cd, c_dist), (name, impl.__name__, "crosscheck")
        md_q, md_t = ref_filter_match_descriptors(k_idx, k_dist, len(t), 64.0, 0.75)
        mf_t, mf_q = ref_filter_match_features(c_idx, c_dist, 64.0)
        np.savez_compressed(os.path.join(GOLD, f"bf_{name}.npz"), q=q, t=t, knn_idx=k_idx, knn_dist=k_dist,
                            cc_idx=c_idx, cc_dist=c_dist, md_query=md_q, md_train=md_t, mf_train=mf_t, mf_query=mf_q,
                            max_dist=np.float32(64.0), ratio=np.float32(0.75))
        print(f"bf_{name}: knn rows {len(q)}, md matches {len(md_q)}, cc matches {int((c_idx >= 0).sum())}, "
              f"mf matches {len(mf_q)}")
        n_bf += 1

    # larger one-off agreement checks (not stored): cv2 vs model vs oracle
    for (nq, nt, seed, gen) in [(2000, 2000, 11, synthetic.bf_correlated), (1500, 1200, 12, synthetic.bf_tie_heavy)]:
        q, t = gen(nq, nt, seed)
        a, b = cv2_knn2(q, t)
        c, d = cpu_oracle.knn2(q, t)
        assert np.array_equal(a, c) and np.array_equal(b, d)
        a, b = cv2_crosscheck(q, t)
        c, d = cpu_oracle.crosscheck(q, t)
        assert np.array_equal(a, c) and np.array_equal(b, d)
        print(f"large check {gen.__name__} {nq}x{nt}: cv2 == oracle (knn2, crosscheck)")

    # KD-tree: the reference's own KDTree2D vs the oracle's implicit-permutation form
    for name, nkp, w, h, seed, radius in [("int_300", 300, 64, 48, 21, 6.0), ("int_2000", 2000, 1241, 376, 22, 20.0),
                                          ("dup_500", 500, 20, 20, 23, 3.0), ("tiny_1", 1, 10, 10, 24, 5.0),
                                          ("tiny_2", 2, 10, 10, 25, 5.0), ("tiny_3", 3, 10, 10, 26, 5.0)]:
        xy = synthetic.keypoints_integer(nkp, w, h, seed)
        rng = np.random.default_rng(seed)
        targets = np.stack([rng.uniform(-2, w + 2, 96), rng.uniform(-2, h + 2, 96)], 1).astype(np.float32)
        targets[:16] = np.rint(targets[:16])            # integer targets: delta == 0 and exact-radius hits
        ref = cpu_oracle.ref_kd_radius_search_all(xy, targets, radius)
        perm = cpu_oracle.kd_build(xy)
        for tgt, r in zip(targets, ref):
            mine = cpu_oracle.kd_radius_search(xy, perm, tgt, radius)
            assert np.array_equal(mine, r), (name, tgt)
        counts = np.array([len(r) for r in ref], np.int32)
        flat = np.concatenate(ref).astype(np.int32) if counts.sum() else np.zeros(0, np.int32)
        np.savez_compressed(os.path.join(GOLD, f"kd_{name}.npz"), xy=xy, targets=targets, radius=np.float32(radius),
                            counts=counts, flat=flat, perm=perm)
        print(f"kd_{name}: {nkp} keypoints, {len(targets)} searches, {int(counts.sum())} hits (reference KDTree2D)")

    # projection scenes (oracle outputs as regression pins)
    for name, npts, nkp, seed, cam, radius, replace in [
            ("small_r20", 600, 300, 31, dict(synthetic.KITTI), 20.0, 0),
            ("small_r15_fuse", 600, 300, 32, dict(synthetic.KITTI), 15.0, 1),
            ("dense_ties", 400, 120, 33, dict(fx=100.0, fy=100.0, cx=40.0, cy=30.0, width=80, height=60), 20.0, 0)]:
        sc = synthetic.projection_scene(npts, nkp, seed, cam=cam)
        if name == "dense_ties":   # low-entropy descriptors -> equal Hamming minima -> KD-order ties
            rng = np.random.default_rng(seed)
            sc.kp_desc[:] = (rng.integers(0, 2, sc.kp_desc.shape, dtype=np.uint8) * 255)
            sc.obs_desc[:] = (rng.integers(0, 2, sc.obs_desc.shape, dtype=np.uint8) * 255)
        perm = cpu_oracle.kd_build(sc.kp_xy)
        prm = cpu_oracle.ProjParams(cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["width"], cam["height"], radius,
                                    64.0 if name != "dense_ties" else 200.0, 0.5, 2.0, 1.25, replace)
        res = cpu_oracle.match_projection(sc.kp_xy, sc.kp_desc, perm, sc.kp_matched, sc.pos, sc.normal, sc.range,
                                          sc.obs_offset, sc.obs_desc, sc.skip, None, sc.pose, prm)
        np.savez_compressed(os.path.join(GOLD, f"proj_{name}.npz"), kp_xy=sc.kp_xy, kp_desc=sc.kp_desc, kd_perm=perm,
                            kp_matched=sc.kp_matched, pos=sc.pos, normal=sc.normal, range=sc.range,
                            obs_offset=sc.obs_offset, obs_desc=sc.obs_desc, skip=sc.skip, pose=sc.pose,
                            cam=np.array([cam["fx"], cam["fy"], cam["cx"], cam["cy"]], np.float32),
                            wh=np.array([cam["width"], cam["height"]], np.int32),
                            params=np.array([radius, prm.max_dist, 0.5, 2.0, 1.25], np.float32),
                            replace=np.int32(replace), out_point=res.point, out_dist=res.dist, uv=res.uv, kp=res.kp,
                            d=res.d)
        print(f"proj_{name}: {npts} points -> {int((res.point >= 0).sum())} matched keypoints, "
              f"{int((res.kp >= 0).sum())} proposing points, near-border {res.near_border}, "
              f"near-radius {res.near_radius}")
    print("golden vectors written to", GOLD)


if __name__ == "__main__":
    main()
