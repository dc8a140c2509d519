"""ctypes view of oracle/liboracle.so (and oracle/_ref/libref_kdtree.so).

TEST INFRASTRUCTURE ONLY -- the checker, never the product.  Importers allowed: tests/,
__graft_entry__.smoke(), bench.py's cpu_baseline / --impl reference legs.  The product package
(racing-slam_b200/) never imports this module.

Reference restated: see the header of oracle/slam_oracle.cpp (file:line citations there).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_REF = None


def build(verbose: bool = False) -> None:
    """Compile liboracle.so (and oracle/_ref when /root/reference exists) via oracle/Makefile."""
    r = subprocess.run(["make", "-C", _HERE], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout, r.stderr)
    if r.returncode != 0:
        raise RuntimeError("oracle build failed")


def _ptr(a, ctype):
    if a is None:
        return None
    return a.ctypes.data_as(C.POINTER(ctype))


class ProjParams(C.Structure):
    """Mirror of OrcProjParams; constants of src/MapMatcher.cpp:12-18 as runtime values."""
    _fields_ = [("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("width", C.c_int32), ("height", C.c_int32), ("radius", C.c_float), ("max_dist", C.c_float),
                ("min_cos", C.c_float), ("nearer", C.c_float), ("further", C.c_float), ("replace", C.c_int32)]


def lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            build()
        L = C.CDLL(path)
        u8p, i32p, f32p, u32p = (C.POINTER(C.c_uint8), C.POINTER(C.c_int32), C.POINTER(C.c_float),
                                 C.POINTER(C.c_uint32))
        L.orc_hamming256.argtypes = [u8p, u8p]
        L.orc_hamming256.restype = C.c_int
        L.orc_knn2.argtypes = [u8p, C.c_int, u8p, C.c_int, i32p, i32p]
        L.orc_crosscheck.argtypes = [u8p, C.c_int, u8p, C.c_int, i32p, i32p]
        L.orc_match_descriptors.argtypes = [u8p, C.c_int, u8p, C.c_int, C.c_float, C.c_float, i32p, i32p]
        L.orc_match_descriptors.restype = C.c_int
        L.orc_match_features.argtypes = [u8p, C.c_int, u8p, C.c_int, C.c_float, i32p, i32p]
        L.orc_match_features.restype = C.c_int
        L.orc_kd_build.argtypes = [f32p, C.c_int, u32p]
        L.orc_kd_radius_search.argtypes = [f32p, u32p, C.c_int, C.c_float, C.c_float, C.c_float, i32p, C.c_int]
        L.orc_kd_radius_search.restype = C.c_int
        L.orc_camera_center.argtypes = [f32p, f32p]
        L.orc_projection_matrix.argtypes = [f32p, C.c_float, C.c_float, C.c_float, C.c_float, f32p]
        L.orc_project.argtypes = [f32p, C.c_float, C.c_float, C.c_float, C.c_float, f32p, C.c_int, f32p]
        L.orc_point_stats.argtypes = [f32p, C.c_int, i32p, f32p, f32p, f32p]
        L.orc_match_projection.argtypes = [C.c_int, f32p, u8p, u32p, u8p, C.c_int, f32p, f32p, f32p, i32p, u8p, u8p,
                                           i32p, C.c_int, f32p, C.POINTER(ProjParams), i32p, i32p, f32p, i32p, i32p,
                                           C.POINTER(C.c_int64)]
        f64p = C.POINTER(C.c_double)
        L.orc_voc_transform_each.argtypes = [i32p, i32p, u8p, f64p, i32p, u8p, C.c_int, i32p, f64p]
        L.orc_voc_bow.argtypes = [i32p, f64p, C.c_int, C.c_int, C.c_int, i32p, f64p]
        L.orc_voc_bow.restype = C.c_int
        L.orc_voc_score_l1.argtypes = [i32p, f64p, C.c_int, i32p, f64p, C.c_int]
        L.orc_voc_score_l1.restype = C.c_double
        L.orc_pnp_project.argtypes = [f64p, f64p, f32p, C.c_int, f32p]
        L.orc_pnp_score.argtypes = [f64p, f64p, f32p, f32p, C.c_int, C.c_double, f32p, u8p]
        L.orc_pnp_score.restype = C.c_int
        L.orc_ransac_update_num_iters.argtypes = [C.c_double, C.c_double, C.c_int, C.c_int]
        L.orc_ransac_update_num_iters.restype = C.c_int
        L.orc_ransac_select.argtypes = [i32p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.POINTER(C.c_int)]
        L.orc_ransac_select.restype = C.c_int
        L.orc_pnp_assemble.argtypes = [i32p, C.c_int, f32p, f32p, f32p, f32p, i32p]
        L.orc_pnp_assemble.restype = C.c_int
        L.orc_pnp_score_batch.argtypes = [f32p, f32p, i32p, i32p, f64p, i32p, C.c_int, f64p, C.c_double, C.c_int, i32p]
        _LIB = L
    return _LIB


def ref_kdtree():
    """The reference's own KDTree2D (oracle/_ref/libref_kdtree.so) or None when not built."""
    global _REF
    if _REF is None:
        path = os.path.join(_HERE, "_ref", "libref_kdtree.so")
        if not os.path.exists(path):
            return None
        R = C.CDLL(path)
        R.ref_kd_create.argtypes = [C.POINTER(C.c_float), C.c_int]
        R.ref_kd_create.restype = C.c_void_p
        R.ref_kd_radius_search.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.POINTER(C.c_int32), C.c_int]
        R.ref_kd_radius_search.restype = C.c_int
        R.ref_kd_destroy.argtypes = [C.c_void_p]
        _REF = R
    return _REF


def _u8(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a.reshape(-1, 32) if a.size else a.reshape(0, 32)


def hamming(a, b) -> int:
    a, b = np.ascontiguousarray(a, np.uint8), np.ascontiguousarray(b, np.uint8)
    return lib().orc_hamming256(_ptr(a, C.c_uint8), _ptr(b, C.c_uint8))


def knn2(q, t):
    q, t = _u8(q), _u8(t)
    idx = np.full((len(q), 2), -1, np.int32)
    dist = np.full((len(q), 2), -1, np.int32)
    if len(q):
        lib().orc_knn2(_ptr(q, C.c_uint8), len(q), _ptr(t, C.c_uint8), len(t), _ptr(idx, C.c_int32),
                       _ptr(dist, C.c_int32))
    return idx, dist


def crosscheck(q, t):
    q, t = _u8(q), _u8(t)
    idx = np.full(len(q), -1, np.int32)
    dist = np.full(len(q), -1, np.int32)
    if len(q):
        lib().orc_crosscheck(_ptr(q, C.c_uint8), len(q), _ptr(t, C.c_uint8), len(t), _ptr(idx, C.c_int32),
                             _ptr(dist, C.c_int32))
    return idx, dist


def match_descriptors(q, t, max_dist=64.0, ratio=0.75):
    """(query_idx, train_idx) arrays; src/MapMatcher.cpp:129-163 with train = matched key-frame rows."""
    q, t = _u8(q), _u8(t)
    oq = np.zeros(max(len(q), 1), np.int32)
    ot = np.zeros(max(len(q), 1), np.int32)
    n = lib().orc_match_descriptors(_ptr(q, C.c_uint8), len(q), _ptr(t, C.c_uint8), len(t), max_dist, ratio,
                                    _ptr(oq, C.c_int32), _ptr(ot, C.c_int32))
    return oq[:n].copy(), ot[:n].copy()


def match_features(prev, cur, max_dist=64.0):
    """(train_idx, query_idx) arrays; src/features/FeatureExtractor.cpp:15-30."""
    prev, cur = _u8(prev), _u8(cur)
    ot = np.zeros(max(len(cur), 1), np.int32)
    oq = np.zeros(max(len(cur), 1), np.int32)
    n = lib().orc_match_features(_ptr(prev, C.c_uint8), len(prev), _ptr(cur, C.c_uint8), len(cur), max_dist,
                                 _ptr(ot, C.c_int32), _ptr(oq, C.c_int32))
    return ot[:n].copy(), oq[:n].copy()


def kd_build(xy):
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    perm = np.zeros(len(xy), np.uint32)
    lib().orc_kd_build(_ptr(xy, C.c_float), len(xy), _ptr(perm, C.c_uint32))
    return perm


def kd_radius_search(xy, perm, target, radius, cap=4096):
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    perm = np.ascontiguousarray(perm, np.uint32)
    out = np.zeros(cap, np.int32)
    n = lib().orc_kd_radius_search(_ptr(xy, C.c_float), _ptr(perm, C.c_uint32), len(xy), float(target[0]),
                                   float(target[1]), float(radius), _ptr(out, C.c_int32), cap)
    assert n <= cap
    return out[:n].copy()


def ref_kd_radius_search_all(xy, targets, radius, cap=4096):
    """Run the REFERENCE KDTree2D on every target; list of index arrays in push_back order."""
    R = ref_kdtree()
    assert R is not None
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    h = R.ref_kd_create(_ptr(xy, C.c_float), len(xy))
    out = np.zeros(cap, np.int32)
    res = []
    try:
        for t in np.asarray(targets, np.float32).reshape(-1, 2):
            n = R.ref_kd_radius_search(h, float(t[0]), float(t[1]), float(radius), _ptr(out, C.c_int32), cap)
            assert n <= cap
            res.append(out[:n].copy())
    finally:
        R.ref_kd_destroy(h)
    return res


def camera_center(pose_colmajor):
    p = np.ascontiguousarray(pose_colmajor, np.float32).reshape(16)
    c = np.zeros(3, np.float32)
    lib().orc_camera_center(_ptr(p, C.c_float), _ptr(c, C.c_float))
    return c


def projection_matrix(pose_colmajor, fx, fy, cx, cy):
    p = np.ascontiguousarray(pose_colmajor, np.float32).reshape(16)
    m = np.zeros(12, np.float32)
    lib().orc_projection_matrix(_ptr(p, C.c_float), fx, fy, cx, cy, _ptr(m, C.c_float))
    return m.reshape(3, 4)


def project(pose_colmajor, fx, fy, cx, cy, pos):
    p = np.ascontiguousarray(pose_colmajor, np.float32).reshape(16)
    pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
    uv = np.zeros((len(pos), 2), np.float32)
    lib().orc_project(_ptr(p, C.c_float), fx, fy, cx, cy, _ptr(pos, C.c_float), len(pos), _ptr(uv, C.c_float))
    return uv


def point_stats(pos, obs_offset, obs_center):
    """avg_viewing_normal / observed_distance_range (src/MapPoint.cpp:24-45) per point."""
    pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
    obs_offset = np.ascontiguousarray(obs_offset, np.int32)
    obs_center = np.ascontiguousarray(obs_center, np.float32).reshape(-1, 3)
    normal = np.zeros((len(pos), 3), np.float32)
    rng = np.zeros((len(pos), 2), np.float32)
    lib().orc_point_stats(_ptr(pos, C.c_float), len(pos), _ptr(obs_offset, C.c_int32), _ptr(obs_center, C.c_float),
                          _ptr(normal, C.c_float), _ptr(rng, C.c_float))
    return normal, rng


@dataclass
class ProjResult:
    point: np.ndarray      # [nkp] point-set index or -1
    dist: np.ndarray       # [nkp] Hamming distance or -1
    uv: np.ndarray         # [n_visit, 2] projections (NaN where the point was skipped up front)
    kp: np.ndarray         # [n_visit] keypoint each visited point chose, -1 = none
    d: np.ndarray          # [n_visit]
    near_border: int
    near_radius: int


def match_projection(kp_xy, kp_desc, kd_perm, kp_matched, pos, normal, rng, obs_offset, obs_desc, skip, order, pose,
                     params: ProjParams, want_stats=True) -> ProjResult:
    kp_xy = np.ascontiguousarray(kp_xy, np.float32).reshape(-1, 2)
    kp_desc = _u8(kp_desc)
    nkp = len(kp_xy)
    kd_perm = np.ascontiguousarray(kd_perm, np.uint32)
    kp_matched = None if kp_matched is None else np.ascontiguousarray(kp_matched, np.uint8)
    pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
    normal = np.ascontiguousarray(normal, np.float32).reshape(-1, 3)
    rng = np.ascontiguousarray(rng, np.float32).reshape(-1, 2)
    obs_offset = np.ascontiguousarray(obs_offset, np.int32)
    obs_desc = _u8(obs_desc)
    skip = None if skip is None else np.ascontiguousarray(skip, np.uint8)
    order = None if order is None else np.ascontiguousarray(order, np.int32)
    n_visit = len(order) if order is not None else len(pos)
    pose = np.ascontiguousarray(pose, np.float32).reshape(16)
    out_point = np.full(nkp, -1, np.int32)
    out_dist = np.full(nkp, -1, np.int32)
    uv = np.zeros((n_visit, 2), np.float32)
    kp = np.zeros(n_visit, np.int32)
    d = np.zeros(n_visit, np.int32)
    stats = (C.c_int64 * 2)(0, 0)
    lib().orc_match_projection(nkp, _ptr(kp_xy, C.c_float), _ptr(kp_desc, C.c_uint8), _ptr(kd_perm, C.c_uint32),
                               _ptr(kp_matched, C.c_uint8), len(pos), _ptr(pos, C.c_float), _ptr(normal, C.c_float),
                               _ptr(rng, C.c_float), _ptr(obs_offset, C.c_int32), _ptr(obs_desc, C.c_uint8),
                               _ptr(skip, C.c_uint8), _ptr(order, C.c_int32), n_visit, _ptr(pose, C.c_float),
                               C.byref(params), _ptr(out_point, C.c_int32), _ptr(out_dist, C.c_int32),
                               _ptr(uv, C.c_float), _ptr(kp, C.c_int32), _ptr(d, C.c_int32),
                               stats if want_stats else None)
    return ProjResult(out_point, out_dist, uv, kp, d, int(stats[0]), int(stats[1]))


# ---------------------------------------------------------------------------------------------
# Bag of words: DBoW2 transform / L1 score, restated in slam_oracle.cpp and pinned
# against the reference's vendored DBoW2 compiled into oracle/_ref/libref_dbow2.so.
# ---------------------------------------------------------------------------------------------
TF_IDF, TF, IDF, BINARY = 0, 1, 2, 3          # DBoW2::WeightingType (BowVector.h:36-42)
NORM_NONE, NORM_L1, NORM_L2 = 0, 1, 2


@dataclass
class Vocabulary:
    """A DBoW2 vocabulary tree as plain arrays (children in DBoW2's own scan order; node 0 = root)."""
    child_offset: np.ndarray    # [n_nodes + 1] i32
    child_ids: np.ndarray       # [n_nodes - 1] i32
    desc: np.ndarray            # [n_nodes, 32] u8
    weight: np.ndarray          # [n_nodes] f64  word weight (idf) at the leaves
    word_id: np.ndarray         # [n_nodes] i32  -1 for inner nodes
    k: int = 0
    L: int = 0


def voc_transform_each(voc: Vocabulary, desc):
    """-> (word [n] i32, weight [n] f64): TemplatedVocabulary::transform(feature, id, weight)."""
    desc = _u8(desc)
    n = len(desc)
    word = np.zeros(n, np.int32)
    w = np.zeros(n, np.float64)
    lib().orc_voc_transform_each(_ptr(voc.child_offset, C.c_int32), _ptr(voc.child_ids, C.c_int32),
                                 _ptr(voc.desc, C.c_uint8), _ptr(voc.weight, C.c_double),
                                 _ptr(voc.word_id, C.c_int32), _ptr(desc, C.c_uint8), n, _ptr(word, C.c_int32),
                                 _ptr(w, C.c_double))
    return word, w


def voc_bow(word, weight, weighting=TF_IDF, norm=NORM_L1):
    """-> (ids [m] i32 ascending, values [m] f64): TemplatedVocabulary::transform(features, bow)."""
    word = np.ascontiguousarray(word, np.int32)
    weight = np.ascontiguousarray(weight, np.float64)
    ids = np.zeros(max(len(word), 1), np.int32)
    vals = np.zeros(max(len(word), 1), np.float64)
    m = lib().orc_voc_bow(_ptr(word, C.c_int32), _ptr(weight, C.c_double), len(word), weighting, norm,
                          _ptr(ids, C.c_int32), _ptr(vals, C.c_double))
    return ids[:m].copy(), vals[:m].copy()


def voc_score_l1(ids_a, vals_a, ids_b, vals_b) -> float:
    ids_a, ids_b = np.ascontiguousarray(ids_a, np.int32), np.ascontiguousarray(ids_b, np.int32)
    vals_a, vals_b = np.ascontiguousarray(vals_a, np.float64), np.ascontiguousarray(vals_b, np.float64)
    return float(lib().orc_voc_score_l1(_ptr(ids_a, C.c_int32), _ptr(vals_a, C.c_double), len(ids_a),
                                        _ptr(ids_b, C.c_int32), _ptr(vals_b, C.c_double), len(ids_b)))


# ---- verify_pnp's input assembly and RANSAC inlier scoring (src/LoopDetector.cpp:176-229) ----
def pnp_project(rt, K4, obj):
    """-> uv [n, 2] f32: cv::projectPoints without distortion; rt = R row-major (9) + t (3), K4 = fx, fy, cx, cy."""
    rt = np.ascontiguousarray(rt, np.float64).reshape(12)
    K4 = np.ascontiguousarray(K4, np.float64).reshape(4)
    obj = np.ascontiguousarray(obj, np.float32).reshape(-1, 3)
    uv = np.zeros((len(obj), 2), np.float32)
    lib().orc_pnp_project(_ptr(rt, C.c_double), _ptr(K4, C.c_double), _ptr(obj, C.c_float), len(obj), _ptr(uv, C.c_float))
    return uv


def pnp_score(rt, K4, obj, img, reproj_error=4.0):
    """-> (count, err [n] f32, mask [n] u8): computeError + findInliers of cv::solvePnPRansac for one model."""
    rt = np.ascontiguousarray(rt, np.float64).reshape(12)
    K4 = np.ascontiguousarray(K4, np.float64).reshape(4)
    obj = np.ascontiguousarray(obj, np.float32).reshape(-1, 3)
    img = np.ascontiguousarray(img, np.float32).reshape(-1, 2)
    err = np.zeros(len(obj), np.float32)
    mask = np.zeros(len(obj), np.uint8)
    c = lib().orc_pnp_score(_ptr(rt, C.c_double), _ptr(K4, C.c_double), _ptr(obj, C.c_float), _ptr(img, C.c_float),
                            len(obj), float(reproj_error), _ptr(err, C.c_float), _ptr(mask, C.c_uint8))
    return int(c), err, mask


def pnp_score_batch(obj, img, corr_begin, corr_count, hyp_rt, hyp_problem, K4, reproj_error=4.0, threads=1):
    """-> inlier count of every model [n_hyp] i32 (ragged problems; OpenMP over models)."""
    obj = np.ascontiguousarray(obj, np.float32)
    img = np.ascontiguousarray(img, np.float32)
    cb, cc = np.ascontiguousarray(corr_begin, np.int32), np.ascontiguousarray(corr_count, np.int32)
    rt, hp = np.ascontiguousarray(hyp_rt, np.float64), np.ascontiguousarray(hyp_problem, np.int32)
    K4 = np.ascontiguousarray(K4, np.float64).reshape(4)
    out = np.zeros(len(hp), np.int32)
    lib().orc_pnp_score_batch(_ptr(obj, C.c_float), _ptr(img, C.c_float), _ptr(cb, C.c_int32), _ptr(cc, C.c_int32),
                              _ptr(rt, C.c_double), _ptr(hp, C.c_int32), len(hp), _ptr(K4, C.c_double),
                              float(reproj_error), int(threads), _ptr(out, C.c_int32))
    return out


def ransac_update_num_iters(p, ep, model_points, max_iters) -> int:
    return int(lib().orc_ransac_update_num_iters(float(p), float(ep), int(model_points), int(max_iters)))


def ransac_select(good_count, n_points, model_points=5, confidence=0.99, max_iters=200):
    """-> (kept model index or -1, models looked at): RANSACPointSetRegistrator::run's choice."""
    g = np.ascontiguousarray(good_count, np.int32)
    it = C.c_int(0)
    b = lib().orc_ransac_select(_ptr(g, C.c_int32), len(g), int(n_points), int(model_points), float(confidence),
                                int(max_iters), C.byref(it))
    return int(b), int(it.value)


def pnp_assemble(match_train, train_pos, query_xy):
    """-> (object [n, 3], image [n, 2], query row [n]): src/LoopDetector.cpp:188-197."""
    mt = np.ascontiguousarray(match_train, np.int32)
    tp = np.ascontiguousarray(train_pos, np.float32).reshape(-1, 3)
    qx = np.ascontiguousarray(query_xy, np.float32).reshape(-1, 2)
    obj = np.zeros((len(mt), 3), np.float32)
    img = np.zeros((len(mt), 2), np.float32)
    qi = np.zeros(len(mt), np.int32)
    n = lib().orc_pnp_assemble(_ptr(mt, C.c_int32), len(mt), _ptr(tp, C.c_float), _ptr(qx, C.c_float),
                               _ptr(obj, C.c_float), _ptr(img, C.c_float), _ptr(qi, C.c_int32))
    return obj[:n].copy(), img[:n].copy(), qi[:n].copy()


_REF_DBOW = None


def ref_dbow2():
    """The reference's own DBoW2 (oracle/_ref/libref_dbow2.so) or None when not built."""
    global _REF_DBOW
    if _REF_DBOW is None:
        path = os.path.join(_HERE, "_ref", "libref_dbow2.so")
        if not os.path.exists(path):
            return None
        R = C.CDLL(path)
        u8p, i32p, f64p = C.POINTER(C.c_uint8), C.POINTER(C.c_int32), C.POINTER(C.c_double)
        R.ref_voc_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, u8p, i32p, C.c_int]
        R.ref_voc_create.restype = C.c_void_p
        R.ref_voc_destroy.argtypes = [C.c_void_p]
        R.ref_voc_num_nodes.argtypes = [C.c_void_p]
        R.ref_voc_num_words.argtypes = [C.c_void_p]
        R.ref_voc_export.argtypes = [C.c_void_p, i32p, i32p, u8p, f64p, i32p]
        R.ref_voc_transform_each.argtypes = [C.c_void_p, u8p, C.c_int, i32p, f64p]
        R.ref_voc_transform_bow.argtypes = [C.c_void_p, u8p, C.c_int, i32p, f64p, C.c_int]
        R.ref_voc_transform_bow.restype = C.c_int
        R.ref_voc_score.argtypes = [C.c_void_p, i32p, f64p, C.c_int, i32p, f64p, C.c_int]
        R.ref_voc_score.restype = C.c_double
        _REF_DBOW = R
    return _REF_DBOW


class RefVocabulary:
    """DBoW2's own TemplatedVocabulary<FORB>: create() from training images, then its transform / score."""

    def __init__(self, k, L, training_desc, image_offset, weighting=TF_IDF, scoring=0):
        self.R = ref_dbow2()
        if self.R is None:
            raise RuntimeError("oracle/_ref/libref_dbow2.so is not built (needs /root/reference)")
        d = _u8(training_desc)
        off = np.ascontiguousarray(image_offset, np.int32)
        self.k, self.L = k, L
        self.h = self.R.ref_voc_create(k, L, weighting, scoring, _ptr(d, C.c_uint8), _ptr(off, C.c_int32), len(off) - 1)

    def export(self) -> Vocabulary:
        n = self.R.ref_voc_num_nodes(self.h)
        v = Vocabulary(np.zeros(n + 1, np.int32), np.zeros(max(n - 1, 1), np.int32), np.zeros((n, 32), np.uint8),
                       np.zeros(n, np.float64), np.zeros(n, np.int32), self.k, self.L)
        self.R.ref_voc_export(self.h, _ptr(v.child_offset, C.c_int32), _ptr(v.child_ids, C.c_int32),
                              _ptr(v.desc, C.c_uint8), _ptr(v.weight, C.c_double), _ptr(v.word_id, C.c_int32))
        v.child_ids = v.child_ids[:n - 1].copy()
        return v

    def transform_each(self, desc):
        desc = _u8(desc)
        word = np.zeros(len(desc), np.int32)
        w = np.zeros(len(desc), np.float64)
        self.R.ref_voc_transform_each(self.h, _ptr(desc, C.c_uint8), len(desc), _ptr(word, C.c_int32),
                                      _ptr(w, C.c_double))
        return word, w

    def transform_bow(self, desc):
        desc = _u8(desc)
        ids = np.zeros(max(len(desc), 1), np.int32)
        vals = np.zeros(max(len(desc), 1), np.float64)
        m = self.R.ref_voc_transform_bow(self.h, _ptr(desc, C.c_uint8), len(desc), _ptr(ids, C.c_int32),
                                         _ptr(vals, C.c_double), len(ids))
        return ids[:m].copy(), vals[:m].copy()

    def score(self, a, b) -> float:
        ia, va = np.ascontiguousarray(a[0], np.int32), np.ascontiguousarray(a[1], np.float64)
        ib, vb = np.ascontiguousarray(b[0], np.int32), np.ascontiguousarray(b[1], np.float64)
        return float(self.R.ref_voc_score(self.h, _ptr(ia, C.c_int32), _ptr(va, C.c_double), len(ia),
                                          _ptr(ib, C.c_int32), _ptr(vb, C.c_double), len(ib)))

    def close(self):
        if self.h:
            self.R.ref_voc_destroy(self.h)
            self.h = None
