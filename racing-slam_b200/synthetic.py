"""Synthetic workloads of the shapes BASELINE.json / SURVEY.md section 8d name (numpy only).

Pure input generators: no matching logic lives here.  Used by tests/, bench.py and smoke().
Shapes follow the reference's data model: 32-byte ORB descriptors (cv::Mat N x 32 CV_8U,
src/features/FeatureExtractor.h:15-18), keypoints as (x, y) float pixels, a 4x4 column-major
pose (Eigen::Matrix4f, src/Frame.h:46), map points with one descriptor per observation
(src/MapPoint.h:38, src/MapMatcher.cpp:86).
"""
from __future__ import annotations

from dataclasses import dataclass

import numpy as np

KITTI = dict(fx=718.856, fy=718.856, cx=607.19, cy=185.22, width=1241, height=376)
HD1080 = dict(fx=1400.0, fy=1400.0, cx=960.0, cy=540.0, width=1920, height=1080)


def flip_bits(desc: np.ndarray, nflip: np.ndarray, rng: np.random.Generator) -> np.ndarray:
    """Return a copy of desc [n,32] u8 with nflip[i] distinct random bits flipped in row i."""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    n = len(desc)
    nflip = np.broadcast_to(np.asarray(nflip, np.int64), (n,))
    bits = np.unpackbits(desc, axis=1)                      # [n,256]
    # rank random keys per row; flip the nflip[i] smallest
    keys = rng.random((n, 256))
    ranks = keys.argsort(axis=1).argsort(axis=1)
    mask = ranks < nflip[:, None]
    bits = bits ^ mask.astype(np.uint8)
    return np.packbits(bits, axis=1)


def bf_uniform(nq: int, nt: int, seed: int):
    rng = np.random.default_rng(seed)
    return (rng.integers(0, 256, (nq, 32), dtype=np.uint8), rng.integers(0, 256, (nt, 32), dtype=np.uint8))


def bf_correlated(nq: int, nt: int, seed: int, max_flip: int = 80, frac_replaced: float = 0.1):
    """train = permuted copy of (part of) query with 0..max_flip flipped bits, so that the
    max-distance 64 and Lowe-ratio gates both pass and fail on real fractions of rows."""
    rng = np.random.default_rng(seed)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    src = rng.integers(0, nq, nt) if nq > 0 else np.zeros(nt, np.int64)
    t = flip_bits(q[src], rng.integers(0, max_flip + 1, nt), rng) if nq > 0 else np.zeros((nt, 32), np.uint8)
    repl = rng.random(nt) < frac_replaced
    t[repl] = rng.integers(0, 256, (int(repl.sum()), 32), dtype=np.uint8)
    perm = rng.permutation(nt)
    return q, np.ascontiguousarray(t[perm])


def bf_tie_heavy(nq: int, nt: int, seed: int, pool: int = 24):
    """Duplicated, low-entropy descriptors: many exact distance ties in rows AND columns, to
    exercise the lowest-index tie-break."""
    rng = np.random.default_rng(seed)
    base = rng.integers(0, 4, (pool, 32), dtype=np.uint8) * 85      # bytes in {0,85,170,255}
    q = base[rng.integers(0, pool, nq)].copy()
    t = base[rng.integers(0, pool, nt)].copy()
    # a sprinkle of single-bit perturbations keeps a few unique minima
    if nq:
        q[rng.integers(0, nq, max(nq // 8, 1)), rng.integers(0, 32)] ^= 1
    if nt:
        t[rng.integers(0, nt, max(nt // 8, 1)), rng.integers(0, 32)] ^= 2
    return q, t


def keypoints_integer(nkp: int, width: int, height: int, seed: int, distinct: bool = False) -> np.ndarray:
    """Keypoints on integer pixels (GFTT corners, src/features/OrbFeatureExtractor.h:25): many
    equal x or y values, which exercises std::nth_element ties in the KD build."""
    rng = np.random.default_rng(seed)
    x = rng.integers(0, width, nkp)
    y = rng.integers(0, height, nkp)
    if distinct:
        cells = rng.choice(width * height, nkp, replace=False)
        x, y = cells % width, cells // width
    return np.stack([x, y], 1).astype(np.float32)


def rot_xyz(rx: float, ry: float, rz: float) -> np.ndarray:
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def make_pose(rng: np.random.Generator, max_deg: float = 2.0, max_t: float = 0.1) -> np.ndarray:
    """World->camera 4x4 (float32, row-major numpy array; flatten(order='F') gives Eigen layout)."""
    ang = np.deg2rad(rng.uniform(-max_deg, max_deg, 3))
    T = np.eye(4)
    T[:3, :3] = rot_xyz(*ang)
    T[:3, 3] = rng.uniform(-max_t, max_t, 3)
    return T.astype(np.float32)


def pose_colmajor(T: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(np.asarray(T, np.float32).flatten(order="F"))


@dataclass
class Scene:
    """One frame + one point set, as POD arrays (the flattening the adapter performs)."""
    cam: dict
    pose: np.ndarray            # [16] column-major world->camera
    kp_xy: np.ndarray           # [nkp,2] f32
    kp_desc: np.ndarray         # [nkp,32] u8
    kp_matched: np.ndarray      # [nkp] u8   Frame::is_matched(size_t), src/Frame.cpp:143-146
    pos: np.ndarray             # [np,3] f32 MapPoint::position()
    normal: np.ndarray          # [np,3] f32 avg_viewing_normal(), src/MapPoint.cpp:24-33
    range: np.ndarray           # [np,2] f32 observed_distance_range(), :35-45
    obs_offset: np.ndarray      # [np+1] i32
    obs_desc: np.ndarray        # [sum_obs,32] u8, host iteration order of observations()
    obs_center: np.ndarray      # [sum_obs,3] f32 camera centres of the observers
    skip: np.ndarray            # [np] u8    is_matched(point) / !is_observed_by(kf)


def point_stats_numpy(pos, obs_offset, obs_center):
    """Inputs for synthetic scenes (float32 numpy; NOT the parity path -- normals and ranges are
    inputs to the matcher, the oracle's orc_point_stats is what restates src/MapPoint.cpp)."""
    n = len(pos)
    counts = np.diff(obs_offset)
    owner = np.repeat(np.arange(n), counts)
    d = (pos[owner] - obs_center).astype(np.float32)
    dist = np.sqrt((d * d).sum(1)).astype(np.float32)
    v = d / np.maximum(dist, 1e-20)[:, None]
    acc = np.zeros((n, 3), np.float32)
    np.add.at(acc, owner, v.astype(np.float32))
    nn = np.sqrt((acc * acc).sum(1))
    normal = (acc / np.maximum(nn, 1e-20)[:, None]).astype(np.float32)
    nearest = np.full(n, np.finfo(np.float32).max, np.float32)
    furthest = np.zeros(n, np.float32)
    np.minimum.at(nearest, owner, dist)
    np.maximum.at(furthest, owner, dist)
    return normal, np.stack([nearest, furthest], 1).astype(np.float32)


def projection_scene(n_points: int, nkp: int, seed: int, cam: dict = KITTI, frac_with_kp: float = 0.3,
                     max_obs: int = 8, obs_flip: int = 40, kp_flip: int = 30, frac_matched_kp: float = 0.1,
                     frac_skip: float = 0.05, depth=(5.0, 50.0), frac_outside: float = 0.1,
                     kp_offset: float = 12.0) -> Scene:
    """points at depth 5-50 in front of a slightly perturbed camera, 1..max_obs
    observations per point (each a <= obs_flip-bit perturbation of a base descriptor), ~30 % of the
    points given a keypoint near their projection, keypoints on integer pixels."""
    rng = np.random.default_rng(seed)
    W, H = cam["width"], cam["height"]
    T = make_pose(rng)
    R, t = T[:3, :3].astype(np.float64), T[:3, 3].astype(np.float64)

    # points: sample pixel + depth in the camera frame, lift to world.  A fraction lands outside
    # the image or behind the camera to exercise P1/P2.
    u = rng.uniform(0, W, n_points)
    v = rng.uniform(0, H, n_points)
    z = rng.uniform(depth[0], depth[1], n_points)
    outside = rng.random(n_points) < frac_outside
    u[outside] += rng.choice([-1.5, 1.5], int(outside.sum())) * W
    behind = rng.random(n_points) < frac_outside * 0.3
    z[behind] *= -1.0
    Xc = np.stack([(u - cam["cx"]) / cam["fx"] * z, (v - cam["cy"]) / cam["fy"] * z, z], 1)
    pos = ((Xc - t) @ R).astype(np.float32)                 # R^T (Xc - t)

    base = rng.integers(0, 256, (n_points, 32), dtype=np.uint8)
    n_obs = np.clip(rng.poisson(3.0, n_points) + 1, 1, max_obs)
    obs_offset = np.zeros(n_points + 1, np.int32)
    obs_offset[1:] = np.cumsum(n_obs)
    owner = np.repeat(np.arange(n_points), n_obs)
    total = int(obs_offset[-1])
    obs_desc = flip_bits(base[owner], rng.integers(0, obs_flip + 1, total), rng)
    # observers: cameras within ~1.5 units of the origin; some far away so that the distance and
    # viewing-angle gates (src/MapMatcher.cpp:62-73) reject a real fraction
    obs_center = rng.normal(0.0, 0.6, (total, 3)).astype(np.float32)
    far = rng.random(total) < 0.05
    obs_center[far] += rng.normal(0.0, 30.0, (int(far.sum()), 3)).astype(np.float32)
    normal, rng_ = point_stats_numpy(pos, obs_offset, obs_center)

    # keypoints on integer pixels: a share sits near point projections with a related descriptor
    kp_xy = keypoints_integer(nkp, W, H, seed + 1000)
    kp_desc = rng.integers(0, 256, (nkp, 32), dtype=np.uint8)
    n_tied = min(int(frac_with_kp * n_points), nkp)
    inside = np.nonzero(~outside & ~behind)[0]
    if n_tied and len(inside):
        pts = rng.choice(inside, n_tied, replace=len(inside) < n_tied)
        kps = rng.choice(nkp, n_tied, replace=False)
        off = rng.uniform(-kp_offset, kp_offset, (n_tied, 2))
        xy = np.rint(np.stack([u[pts], v[pts]], 1) + off)
        xy[:, 0] = np.clip(xy[:, 0], 0, W - 1)
        xy[:, 1] = np.clip(xy[:, 1], 0, H - 1)
        kp_xy[kps] = xy.astype(np.float32)
        kp_desc[kps] = flip_bits(base[pts], rng.integers(0, kp_flip + 1, n_tied), rng)
    kp_matched = (rng.random(nkp) < frac_matched_kp).astype(np.uint8)
    skip = (rng.random(n_points) < frac_skip).astype(np.uint8)
    return Scene(cam=dict(cam), pose=pose_colmajor(T), kp_xy=kp_xy, kp_desc=kp_desc, kp_matched=kp_matched, pos=pos,
                 normal=normal, range=rng_, obs_offset=obs_offset, obs_desc=obs_desc, obs_center=obs_center, skip=skip)


def next_frame_descriptors(desc: np.ndarray, seed: int, max_flip: int = 30, frac_replaced: float = 0.1):
    """descriptors of frame f+1 = frame f with <= 30 bits flipped, 10 % replaced,
    rows shuffled."""
    rng = np.random.default_rng(seed)
    n = len(desc)
    out = flip_bits(desc, rng.integers(0, max_flip + 1, n), rng)
    repl = rng.random(n) < frac_replaced
    out[repl] = rng.integers(0, 256, (int(repl.sum()), 32), dtype=np.uint8)
    return np.ascontiguousarray(out[rng.permutation(n)])


def keyframe_db(n_frames: int, n_desc: int, seed: int, shared_frac: float = 0.3, max_flip: int = 40) -> np.ndarray:
    """n_frames x n_desc x 32 key-frame descriptor database.  A share of every
    frame's descriptors are perturbed copies from a common landmark pool, so the max-distance and
    ratio gates pass for a realistic fraction of rows."""
    rng = np.random.default_rng(seed)
    pool = rng.integers(0, 256, (max(n_desc, 1), 32), dtype=np.uint8)
    db = rng.integers(0, 256, (n_frames, n_desc, 32), dtype=np.uint8)
    n_shared = int(shared_frac * n_desc)
    for f in range(n_frames):
        if n_shared == 0:
            break
        rows = rng.choice(n_desc, n_shared, replace=False)
        src = rng.choice(len(pool), n_shared, replace=False)
        db[f, rows] = flip_bits(pool[src], rng.integers(0, max_flip + 1, n_shared), rng)
    return db


def keyframe_db_fast(n_frames: int, n_desc: int, seed: int, shared_frac: float = 0.3) -> np.ndarray:
    """Same shape and intent as keyframe_db, cheap enough for the full C5 size (1000 x 4000 x 32 B
    = 128 MB): shared rows are pool rows XOR a sparse mask (each bit set with probability 1/8,
    about 32 of 256 bits), so they pass the distance-64 and ratio gates; the rest is uniform."""
    rng = np.random.default_rng(seed)
    db = rng.integers(0, 256, (n_frames, n_desc, 32), dtype=np.uint8)
    pool = rng.integers(0, 256, (max(n_desc, 1), 32), dtype=np.uint8)
    n_shared = int(shared_frac * n_desc)
    if n_shared == 0:
        return db
    for f in range(n_frames):
        rows = rng.choice(n_desc, n_shared, replace=False)
        src = rng.choice(len(pool), n_shared, replace=False)
        m = rng.integers(0, 256, (3, n_shared, 32), dtype=np.uint8)
        db[f, rows] = pool[src] ^ (m[0] & m[1] & m[2])
    return db


class TrackingSequence:
    """a synthetic tracking run.  One pool of map points laid out for a base camera
    pose; the map at frame f is the pool's first f * points_per_frame points (Map's vector order,
    src/Map.cpp:164-190), so it grows 0 -> n_frames * points_per_frame.  Every frame has the base
    frame's keypoints jittered by <= 1 px (integer pixels) and descriptors that are a fresh <= 30-bit
    perturbation of the base descriptors with 10 % replaced, so consecutive frames differ by <= 60
    bits on the surviving rows and the link to the map descriptors does not drift over the run."""

    def __init__(self, n_frames: int, nkp: int, points_per_frame: int, seed: int, cam: dict = HD1080,
                 kf_window: int = 3000, frac_tracked: float = 0.05):
        self.n_frames, self.nkp, self.ppf, self.seed, self.cam = n_frames, nkp, points_per_frame, seed, dict(cam)
        self.kf_window, self.frac_tracked = kf_window, frac_tracked
        self.pool = projection_scene(n_frames * points_per_frame, nkp, seed, cam=cam)
        self.base_T = np.asarray(self.pool.pose, np.float32).reshape(4, 4, order="F").copy()

    def map_size(self, f: int) -> int:
        """Points in the map when frame f is tracked (the frame's own new points come after it)."""
        return f * self.ppf

    def frame(self, f: int):
        """-> dict(kp_xy, kp_desc, kp_matched, pose, matched_points, kf_points) for frame f."""
        rng = np.random.default_rng(self.seed * 1000003 + f)
        W, H = self.cam["width"], self.cam["height"]
        xy = self.pool.kp_xy + rng.integers(-1, 2, self.pool.kp_xy.shape).astype(np.float32)
        xy[:, 0] = np.clip(xy[:, 0], 0, W - 1)
        xy[:, 1] = np.clip(xy[:, 1], 0, H - 1)
        desc = flip_bits(self.pool.kp_desc, rng.integers(0, 31, self.nkp), rng)
        repl = rng.random(self.nkp) < 0.1
        desc[repl] = rng.integers(0, 256, (int(repl.sum()), 32), dtype=np.uint8)
        T = self.base_T.copy()
        T[:3, 3] += rng.uniform(-0.02, 0.02, 3).astype(np.float32)
        n = self.map_size(f)
        kf_points = np.arange(max(0, n - self.kf_window), n, dtype=np.int32)    # seen by the last key frame
        tracked = kf_points[rng.random(len(kf_points)) < self.frac_tracked]      # track_from_last_frame
        return dict(kp_xy=np.ascontiguousarray(xy, np.float32), kp_desc=np.ascontiguousarray(desc),
                    kp_matched=self.pool.kp_matched, pose=pose_colmajor(T), matched_points=tracked.astype(np.int32),
                    kf_points=kf_points)

    def new_points(self, f: int):
        """The block Mapper::insert adds after frame f: (pos, normal, range, obs_offset (relative), obs_desc)."""
        a, b = f * self.ppf, (f + 1) * self.ppf
        p = self.pool
        o0, o1 = int(p.obs_offset[a]), int(p.obs_offset[b])
        return (p.pos[a:b], p.normal[a:b], p.range[a:b], (p.obs_offset[a:b + 1] - o0).astype(np.int32),
                p.obs_desc[o0:o1])


def random_vocabulary(k: int, L: int, seed: int, flip: int = 48):
    """A synthetic DBoW2-shaped vocabulary tree for sizes DBoW2's own k-means is too slow to
    build here: complete k-ary tree of depth L, node ids in breadth-first order (children of a node are
    contiguous, as DBoW2's HKmeansStep creates them), every child descriptor = its parent's with <= flip
    bits flipped, leaf weights = a plausible idf in (0, 10].  -> dict of the arrays the C ABI takes
    (child_offset, child_ids, desc, weight, word_id).  The ORB vocabulary the reference loads has k = 10,
    L = 6 (about 1.1 M nodes)."""
    rng = np.random.default_rng(seed)
    level_sizes = [k ** l for l in range(L + 1)]
    n = int(sum(level_sizes))
    desc = np.zeros((n, 32), np.uint8)
    child_offset = np.zeros(n + 1, np.int32)
    first = np.cumsum([0] + level_sizes)              # first node id of every level
    for l in range(L):
        parents = np.arange(first[l], first[l + 1])
        kids = np.arange(first[l + 1], first[l + 2])
        base = desc[np.repeat(parents, k)] if l > 0 else rng.integers(0, 256, (len(kids), 32), dtype=np.uint8)
        desc[kids] = flip_bits(base, rng.integers(flip // 2, flip + 1, len(kids)), rng) if l > 0 else base
        child_offset[parents + 1] = k
    child_offset = np.cumsum(child_offset).astype(np.int32)
    child_ids = np.arange(1, n, dtype=np.int32)
    word_id = np.full(n, -1, np.int32)
    leaves = np.arange(first[L], n)
    word_id[leaves] = np.arange(len(leaves), dtype=np.int32)
    weight = np.zeros(n, np.float64)
    weight[leaves] = rng.uniform(0.05, 10.0, len(leaves))
    return dict(child_offset=child_offset, child_ids=child_ids, desc=desc, weight=weight, word_id=word_id, k=k, L=L)


def vocabulary_queries(voc: dict, n: int, seed: int, flip: int = 40) -> np.ndarray:
    """n query descriptors: perturbed copies of random leaf descriptors, so the descent is decisive at most
    levels but not all."""
    rng = np.random.default_rng(seed)
    leaves = np.nonzero(voc["word_id"] >= 0)[0]
    src = voc["desc"][rng.choice(leaves, n)]
    return flip_bits(src, rng.integers(0, flip + 1, n), rng)


# ---- verify_pnp problems (src/LoopDetector.cpp:176-229) ----------------------------------------
PNP_K4 = np.array([KITTI["fx"], KITTI["fy"], KITTI["cx"], KITTI["cy"]], np.float64)


def pnp_problems(seed: int, sizes, n_models, noise: float = 1.0, outliers: float = 0.4):
    """Synthetic verify_pnp problems: per problem n 3-D points in front of a camera with a random true pose, their
    projections (+ Gaussian pixel noise, a fraction replaced by uniform outliers) and h pose models in iteration
    order as [h, 12] doubles (R row-major, t): model 3, 10, 17 ... are the exact pose, the others are perturbed by
    0 .. 4 steps (rotation entries 1.5e-3 per step, translation 2e-2 per step).
    -> (objects [n, 3] f32 list, images [n, 2] f32 list, models [h, 12] f64 list)."""
    rng = np.random.default_rng(seed)
    objs, imgs, models = [], [], []
    for n, h in zip(sizes, n_models):
        obj = (rng.uniform(-8, 8, (n, 3)) + [0, 0, 25]).astype(np.float32)
        R = rot_xyz(*rng.uniform(-0.2, 0.2, 3)).astype(np.float64)
        t = rng.uniform(-0.5, 0.5, 3)
        P = obj.astype(np.float64) @ R.T + t
        img = np.stack([P[:, 0] / P[:, 2] * PNP_K4[0] + PNP_K4[2], P[:, 1] / P[:, 2] * PNP_K4[1] + PNP_K4[3]], 1)
        img = (img + rng.normal(0, noise, img.shape)).astype(np.float32) if noise > 0 else img.astype(np.float32)
        bad = rng.random(n) < outliers
        img[bad] = rng.uniform(0, 1241, (int(bad.sum()), 2)).astype(np.float32)
        step = (np.arange(h) + 2) % 5
        step = np.where(np.arange(h) % 7 == 3, 0, step).astype(np.float64)[:, None]
        rt = np.empty((h, 12), np.float64)
        rt[:, :9] = R.reshape(1, 9) + rng.normal(0, 1.0, (h, 9)) * (1.5e-3 * step)
        rt[:, 9:] = t[None, :] + rng.normal(0, 1.0, (h, 3)) * (2e-2 * step)
        objs.append(obj)
        imgs.append(img)
        models.append(rt)
    return objs, imgs, models
