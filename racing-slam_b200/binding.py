"""ctypes binding of libslamgpu.so (include/slamgpu.h) -- the only way this package computes.

There is NO CPU path here: if the CUDA library is missing or no sm_100 device is present the
calls raise.  Host arrays are numpy; the *_dev forms take raw device pointers (e.g. torch
`tensor.data_ptr()`), enqueue on the context's stream and do not synchronise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libslamgpu.so")
# tools/prof_tc.py only: the timing-experiment build (make -C csrc debug) in place of the product library
if os.environ.get("SLAMGPU_DEBUG_LIBRARY"):
    LIB_PATH = os.environ["SLAMGPU_DEBUG_LIBRARY"]

OK, ERR_INVALID, ERR_CUDA, ERR_NOMEM, ERR_NOTFOUND = 0, 1, 2, 3, 4
MAX_TRAIN = 1 << 22

_u8p, _i32p, _u32p, _f32p = (C.POINTER(C.c_uint8), C.POINTER(C.c_int32), C.POINTER(C.c_uint32), C.POINTER(C.c_float))
_i64p = C.POINTER(C.c_int64)
_f64p = C.POINTER(C.c_double)
BOW_TF_IDF, BOW_TF, BOW_IDF, BOW_BINARY = 0, 1, 2, 3          # DBoW2::WeightingType
BOW_NORM_NONE, BOW_NORM_L1, BOW_NORM_L2 = 0, 1, 2
BOW_MAX_FEATURES = 16384
_ctx = C.c_void_p


class SlamGpuError(RuntimeError):
    def __init__(self, code: int, text: str):
        super().__init__(f"libslamgpu error {code}: {text}")
        self.code = code


class ProjParams(C.Structure):
    """slamgpu_proj_params: constants of src/MapMatcher.cpp:12-18 and the camera, as runtime values."""
    _fields_ = [("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("width", C.c_int32), ("height", C.c_int32), ("radius", C.c_float), ("max_dist", C.c_float),
                ("min_cos", C.c_float), ("nearer", C.c_float), ("further", C.c_float), ("replace", C.c_int32)]


# name -> (restype, argtypes); every symbol include/slamgpu.h declares
SIGNATURES = {
    "slamgpu_abi_version": (C.c_int, []),
    "slamgpu_last_error": (C.c_char_p, []),
    "slamgpu_device_count": (C.c_int, [C.POINTER(C.c_int)]),
    "slamgpu_create": (C.c_int, [C.c_int, C.POINTER(_ctx)]),
    "slamgpu_destroy": (C.c_int, [_ctx]),
    "slamgpu_sync": (C.c_int, [_ctx]),
    "slamgpu_stream": (C.c_void_p, [_ctx]),
    "slamgpu_launch_count": (C.c_int64, [_ctx]),
    "slamgpu_host_alloc": (C.c_int, [C.POINTER(C.c_void_p), C.c_size_t]),
    "slamgpu_host_free": (C.c_int, [C.c_void_p]),
    "slamgpu_knn2": (C.c_int, [_ctx, _u8p, C.c_int, _u8p, C.c_int, _i32p, _i32p]),
    "slamgpu_crosscheck": (C.c_int, [_ctx, _u8p, C.c_int, _u8p, C.c_int, _i32p, _i32p]),
    "slamgpu_match_descriptors": (C.c_int, [_ctx, _u8p, C.c_int, _u8p, C.c_int, C.c_float, C.c_float, _i32p, _i32p,
                                            C.POINTER(C.c_int)]),
    "slamgpu_match_features": (C.c_int, [_ctx, _u8p, C.c_int, _u8p, C.c_int, C.c_float, _i32p, _i32p,
                                         C.POINTER(C.c_int)]),
    "slamgpu_knn2_dev": (C.c_int, [_ctx, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "slamgpu_crosscheck_dev": (C.c_int, [_ctx, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "slamgpu_knn2_dev_noalloc": (C.c_int, [_ctx, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "slamgpu_match_descriptors_dev": (C.c_int, [_ctx, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_float, C.c_float,
                                                C.c_void_p, C.c_void_p]),
    "slamgpu_match_descriptors_batch_dev": (C.c_int, [_ctx, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int,
                                                      C.c_int, C.c_float, C.c_float, C.c_void_p, C.c_void_p]),
    "slamgpu_match_descriptors_batch": (C.c_int, [_ctx, _u8p, _i32p, C.c_int, _i32p, C.c_int, C.c_float, C.c_float,
                                                  _i32p, _i32p, _i32p, C.c_int64, _i64p]),
    "slamgpu_kfdb_append": (C.c_int, [_ctx, C.c_int, _u8p, C.c_int, C.POINTER(C.c_int)]),
    "slamgpu_kfdb_count": (C.c_int, [_ctx, C.c_int, C.POINTER(C.c_int), _i64p]),
    "slamgpu_kfdb_match_pairs": (C.c_int, [_ctx, C.c_int, _i32p, C.c_int, C.c_float, C.c_float, _i32p, _i32p, _i32p,
                                           C.c_int64, _i64p]),
    "slamgpu_kfdb_release": (C.c_int, [_ctx, C.c_int]),
    "slamgpu_frame_upload": (C.c_int, [_ctx, C.c_int, C.c_int, _f32p, _u8p, _u32p, _u8p]),
    "slamgpu_frame_set_matched": (C.c_int, [_ctx, C.c_int, _u8p]),
    "slamgpu_frame_release": (C.c_int, [_ctx, C.c_int]),
    "slamgpu_frame_get_kd_perm": (C.c_int, [_ctx, C.c_int, _u32p]),
    "slamgpu_kd_build_host": (C.c_int, [_f32p, C.c_int, _u32p]),
    "slamgpu_points_upload": (C.c_int, [_ctx, C.c_int, C.c_int, _f32p, _f32p, _f32p, _i32p, _u8p, _f32p]),
    "slamgpu_points_append": (C.c_int, [_ctx, C.c_int, C.c_int, _f32p, _f32p, _f32p, _i32p, _u8p, _f32p]),
    "slamgpu_points_count": (C.c_int, [_ctx, C.c_int, C.POINTER(C.c_int), _i64p]),
    "slamgpu_points_release": (C.c_int, [_ctx, C.c_int]),
    "slamgpu_points_get_stats": (C.c_int, [_ctx, C.c_int, _f32p, _f32p]),
    "slamgpu_points_create": (C.c_int, [_ctx, C.c_int, C.c_int, C.c_int64, C.c_int]),
    "slamgpu_observer_centers_update": (C.c_int, [_ctx, C.c_int, _i32p, C.c_int, _f32p]),
    "slamgpu_points_append_observed": (C.c_int, [_ctx, C.c_int, C.c_int, _f32p, _i32p, _u8p, _i32p]),
    "slamgpu_points_set_observations": (C.c_int, [_ctx, C.c_int, _i32p, C.c_int, _i32p, _u8p, _i32p]),
    "slamgpu_points_set_position": (C.c_int, [_ctx, C.c_int, _i32p, C.c_int, _f32p]),
    "slamgpu_points_remove": (C.c_int, [_ctx, C.c_int, _i32p, C.c_int]),
    "slamgpu_points_compact": (C.c_int, [_ctx, C.c_int, _i32p, C.POINTER(C.c_int)]),
    "slamgpu_points_live": (C.c_int, [_ctx, C.c_int, C.POINTER(C.c_int), _i64p, _i64p]),
    "slamgpu_match_projection": (C.c_int, [_ctx, C.c_int, C.c_int, _f32p, C.POINTER(ProjParams), _u8p, _i32p, C.c_int,
                                           _i32p, _i32p]),
    "slamgpu_match_projection_lists": (C.c_int, [_ctx, C.c_int, C.c_int, _f32p, C.POINTER(ProjParams), _i32p, C.c_int,
                                                 _i32p, C.c_int, _i32p, _i32p]),
    "slamgpu_match_projection_ex": (C.c_int, [_ctx, C.c_int, C.c_int, _f32p, C.POINTER(ProjParams), _u8p, _i32p,
                                              C.c_int, _i32p, _i32p, _f32p, _i32p, _i32p, _i64p]),
    "slamgpu_track_match": (C.c_int, [_ctx, C.c_int, C.c_int, _f32p, C.POINTER(ProjParams), _i32p, C.c_int, _i32p,
                                      C.c_int, _i32p, _i32p, _i32p, _i32p]),
    "slamgpu_match_projection_batch": (C.c_int, [_ctx, C.POINTER(C.c_int), C.c_int, C.c_int, _f32p,
                                                 C.POINTER(ProjParams), _u8p, _i32p, C.c_int, _i32p, _i32p]),
    "slamgpu_vocab_upload": (C.c_int, [_ctx, C.c_int, C.c_int, _i32p, _i32p, _u8p, _f64p, _i32p]),
    "slamgpu_vocab_release": (C.c_int, [_ctx, C.c_int]),
    "slamgpu_vocab_transform": (C.c_int, [_ctx, C.c_int, _u8p, C.c_int, _i32p, _f64p]),
    "slamgpu_vocab_transform_dev": (C.c_int, [_ctx, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    "slamgpu_vocab_bow": (C.c_int, [_ctx, C.c_int, _u8p, _i32p, C.c_int, C.c_int, C.c_int, _i32p, _f64p, _i32p]),
    "slamgpu_vocab_score_l1": (C.c_int, [_ctx, _i32p, _f64p, C.c_int, _i32p, _f64p, _i32p, C.c_int, _f64p]),
    "slamgpu_bowdb_append": (C.c_int, [_ctx, C.c_int, _i32p, _f64p, C.c_int]),
    "slamgpu_bowdb_count": (C.c_int, [_ctx, C.c_int, C.POINTER(C.c_int)]),
    "slamgpu_bowdb_score_l1": (C.c_int, [_ctx, C.c_int, _i32p, _f64p, C.c_int, _f64p]),
    "slamgpu_bowdb_release": (C.c_int, [_ctx, C.c_int]),
    "slamgpu_pnp_assemble_batch_dev": (C.c_int, [_ctx, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                                 C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                 C.c_void_p]),
    "slamgpu_pnp_score_dev": (C.c_int, [_ctx, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                        C.c_void_p, C.c_int, _f64p, C.c_double, C.c_void_p]),
    "slamgpu_pnp_mask_dev": (C.c_int, [_ctx, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                       C.c_void_p, _f64p, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p]),
    "slamgpu_ransac_select": (C.c_int, [_i32p, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.POINTER(C.c_int),
                                        C.POINTER(C.c_int)]),
    "slamgpu_pnp_ransac_score": (C.c_int, [_ctx, _f32p, _f32p, _i32p, C.c_int, _f64p, _i32p, _f64p, C.c_double,
                                           C.c_double, C.c_int, C.c_int, _i32p, _i32p, _i32p, _u8p]),
    "slamgpu_host_project": (C.c_int, [_f32p, C.POINTER(ProjParams), _f32p, _f32p]),
    "slamgpu_host_camera_center": (C.c_int, [_f32p, _f32p]),
    "slamgpu_host_point_stats": (C.c_int, [_f32p, C.c_int, _f32p, _f32p, _f32p]),
    "slamgpu_speculation_count": (C.c_int, [_ctx, _i64p, _i64p]),
    "slamgpu_trace_read": (C.c_int, [_ctx, _f64p, _i64p]),
    "slamgpu_profile_enable": (C.c_int, [_ctx, C.c_int]),
    "slamgpu_profile_read": (C.c_int, [_ctx, C.c_int, C.POINTER(C.c_double), _i64p]),
    "slamgpu_bench_popc": (C.c_int, [_ctx, C.POINTER(C.c_double), C.POINTER(C.c_int), C.POINTER(C.c_double)]),
}

_LIB = None


def load():
    """dlopen the in-tree CUDA library and bind every symbol.  Raises if it has not been built."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(
                f"{LIB_PATH} is missing: build it with `make -C racing-slam_b200/csrc` "
                "(or __graft_entry__.build()).  This package has no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)           # AttributeError = header/library mismatch
            fn.restype = res
            fn.argtypes = args
        _LIB = lib
    return _LIB


def _p(a, ctype):
    return None if a is None else a.ctypes.data_as(C.POINTER(ctype))


def _desc(a):
    a = np.ascontiguousarray(a, dtype=np.uint8)
    return a.reshape(-1, 32)


def _opt(a, dtype):
    return None if a is None else np.ascontiguousarray(a, dtype=dtype)


def kd_build_host(xy) -> np.ndarray:
    """KDTree2D::build as a permutation (src/KDTree.cpp:8-43); host-only, needs no device."""
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    perm = np.zeros(len(xy), np.uint32)
    lib = load()
    rc = lib.slamgpu_kd_build_host(_p(xy, C.c_float), len(xy), _p(perm, C.c_uint32))
    if rc != OK:
        raise SlamGpuError(rc, lib.slamgpu_last_error().decode())
    return perm


def host_project(pose, params: ProjParams, xyz) -> np.ndarray:
    """Camera::project in the library's operation order, on the host (no device)."""
    pose = np.ascontiguousarray(pose, np.float32).reshape(16)
    xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
    out = np.zeros((len(xyz), 2), np.float32)
    lib = load()
    for i in range(len(xyz)):
        lib.slamgpu_host_project(_p(pose, C.c_float), C.byref(params), _p(xyz[i], C.c_float), _p(out[i], C.c_float))
    return out


def host_camera_center(pose) -> np.ndarray:
    pose = np.ascontiguousarray(pose, np.float32).reshape(16)
    out = np.zeros(3, np.float32)
    load().slamgpu_host_camera_center(_p(pose, C.c_float), _p(out, C.c_float))
    return out


def host_point_stats(xyz, obs_center):
    xyz = np.ascontiguousarray(xyz, np.float32).reshape(3)
    obs_center = np.ascontiguousarray(obs_center, np.float32).reshape(-1, 3)
    normal, rng = np.zeros(3, np.float32), np.zeros(2, np.float32)
    load().slamgpu_host_point_stats(_p(xyz, C.c_float), len(obs_center), _p(obs_center, C.c_float),
                                    _p(normal, C.c_float), _p(rng, C.c_float))
    return normal, rng


def ransac_select(good_count, n_points: int, model_points: int = 5, confidence: float = 0.99, max_iters: int = 200):
    """RANSACPointSetRegistrator::run's choice over models in iteration order -> (kept model or -1, models looked at).
    Host arithmetic inside the library; needs no device."""
    g = np.ascontiguousarray(good_count, np.int32)
    best, iters = C.c_int(-1), C.c_int(0)
    rc = load().slamgpu_ransac_select(_p(g, C.c_int32), len(g), int(n_points), int(model_points), float(confidence),
                                      int(max_iters), C.byref(best), C.byref(iters))
    if rc != 0:
        raise SlamGpuError(rc, load().slamgpu_last_error().decode())
    return best.value, iters.value


SLAMGPU_MAX_ROWS_DEFAULT = 4096     # default per-pair capacity of kfdb_match_pairs when the caller gives none


def pinned_empty(shape, dtype) -> np.ndarray:
    """A numpy array over page-locked host memory from slamgpu_host_alloc (freed when the array and its views are
    collected).  Descriptor sets and result arrays handed to the host-buffer calls from such arrays are copied by
    DMA directly, without the driver's staging copy."""
    import weakref
    dtype = np.dtype(dtype)
    n = int(np.prod(shape))
    lib = load()
    ptr = C.c_void_p()
    rc = lib.slamgpu_host_alloc(C.byref(ptr), max(n * dtype.itemsize, 1))
    if rc != OK:
        raise SlamGpuError(rc, lib.slamgpu_last_error().decode())
    buf = (C.c_uint8 * max(n * dtype.itemsize, 1)).from_address(ptr.value)
    weakref.finalize(buf, lib.slamgpu_host_free, C.c_void_p(ptr.value))
    return np.frombuffer(buf, dtype=dtype, count=n).reshape(shape)


def device_count() -> int:
    lib = load()
    n = C.c_int(0)
    lib.slamgpu_device_count(C.byref(n))
    return n.value


class Context:
    """One CUDA device + one stream + pinned staging (slamgpu_ctx)."""

    def __init__(self, device: int = 0):
        self._lib = load()
        self._h = _ctx()
        self._check(self._lib.slamgpu_create(device, C.byref(self._h)))
        self.device = device

    def _check(self, rc: int):
        if rc != OK:
            raise SlamGpuError(rc, self._lib.slamgpu_last_error().decode())

    def close(self):
        if self._h:
            self._lib.slamgpu_destroy(self._h)
            self._h = _ctx()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- plumbing ----
    def sync(self):
        self._check(self._lib.slamgpu_sync(self._h))

    @property
    def stream(self) -> int:
        return int(self._lib.slamgpu_stream(self._h) or 0)

    @property
    def launches(self) -> int:
        return int(self._lib.slamgpu_launch_count(self._h))

    def speculation_count(self):
        """(speculative projection calls, of which repeated with the KD-tree because of a tie)"""
        a, b = C.c_int64(0), C.c_int64(0)
        self._check(self._lib.slamgpu_speculation_count(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def trace_read(self):
        """SLAMGPU_TRACE_BATCH=1: (microseconds [queue, wait, tail, call] summed since the last read, calls)"""
        us = np.zeros(4, np.float64)
        n = C.c_int64(0)
        self._check(self._lib.slamgpu_trace_read(self._h, _p(us, C.c_double), C.byref(n)))
        return us, n.value

    def profile_enable(self, on: bool = True):
        self._check(self._lib.slamgpu_profile_enable(self._h, 1 if on else 0))

    def profile_read(self, kernel: int):
        """(total_ms, launches) of kernel 0 = brute-force scan, 1 = projection match since the last read."""
        ms, n = C.c_double(0), C.c_int64(0)
        self._check(self._lib.slamgpu_profile_read(self._h, kernel, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def bench_popc(self):
        rate, n_sm, mhz = C.c_double(0), C.c_int(0), C.c_double(0)
        self._check(self._lib.slamgpu_bench_popc(self._h, C.byref(rate), C.byref(n_sm), C.byref(mhz)))
        return rate.value, n_sm.value, mhz.value

    # ---- brute force, host buffers ----
    def knn2(self, q, t):
        q, t = _desc(q), _desc(t)
        idx = np.full((len(q), 2), -1, np.int32)
        dist = np.full((len(q), 2), -1, np.int32)
        self._check(self._lib.slamgpu_knn2(self._h, _p(q, C.c_uint8), len(q), _p(t, C.c_uint8), len(t),
                                           _p(idx, C.c_int32), _p(dist, C.c_int32)))
        return idx, dist

    def crosscheck(self, q, t):
        q, t = _desc(q), _desc(t)
        idx = np.full(len(q), -1, np.int32)
        dist = np.full(len(q), -1, np.int32)
        self._check(self._lib.slamgpu_crosscheck(self._h, _p(q, C.c_uint8), len(q), _p(t, C.c_uint8), len(t),
                                                 _p(idx, C.c_int32), _p(dist, C.c_int32)))
        return idx, dist

    def match_descriptors(self, q, t, max_dist=64.0, ratio=0.75):
        """-> (query_idx, train_idx), ascending query order (src/MapMatcher.cpp:129-163)."""
        q, t = _desc(q), _desc(t)
        oq = np.zeros(max(len(q), 1), np.int32)
        ot = np.zeros(max(len(q), 1), np.int32)
        n = C.c_int(0)
        self._check(self._lib.slamgpu_match_descriptors(self._h, _p(q, C.c_uint8), len(q), _p(t, C.c_uint8), len(t),
                                                        max_dist, ratio, _p(oq, C.c_int32), _p(ot, C.c_int32),
                                                        C.byref(n)))
        return oq[:n.value].copy(), ot[:n.value].copy()

    def match_features(self, prev, cur, max_dist=64.0):
        """-> (train_idx, query_idx) (src/features/FeatureExtractor.cpp:15-30)."""
        prev, cur = _desc(prev), _desc(cur)
        ot = np.zeros(max(len(cur), 1), np.int32)
        oq = np.zeros(max(len(cur), 1), np.int32)
        n = C.c_int(0)
        self._check(self._lib.slamgpu_match_features(self._h, _p(prev, C.c_uint8), len(prev), _p(cur, C.c_uint8),
                                                     len(cur), max_dist, _p(ot, C.c_int32), _p(oq, C.c_int32),
                                                     C.byref(n)))
        return ot[:n.value].copy(), oq[:n.value].copy()

    def match_descriptors_batch(self, desc, frame_offset, pairs, max_dist=64.0, ratio=0.75, capacity=None, out=None):
        """-> (offset [n_pairs+1], query_idx, train_idx): per pair (query_frame, train_frame) the
        compacted match_descriptors list.  out = (query_idx, train_idx) int32 arrays to fill (pinned ones receive
        the device copies directly); their length is then the capacity and the returned lists are views of them."""
        desc = _desc(desc)
        frame_offset = np.ascontiguousarray(frame_offset, np.int32)
        pairs = np.ascontiguousarray(pairs, np.int32).reshape(-1, 2)
        n_pairs = len(pairs)
        if capacity is None:   # every query row can match at most once per pair
            rows = np.diff(frame_offset)
            capacity = int(rows[pairs[:, 0]].sum()) if n_pairs else 0
        off = np.zeros(n_pairs + 1, np.int32)
        if out is not None:
            oq, ot = out
            assert oq.dtype == np.int32 and ot.dtype == np.int32 and oq.flags.c_contiguous and ot.flags.c_contiguous
            capacity = min(len(oq), len(ot))
        else:
            oq = np.zeros(max(capacity, 1), np.int32)
            ot = np.zeros(max(capacity, 1), np.int32)
        total = C.c_int64(0)
        self._check(self._lib.slamgpu_match_descriptors_batch(
            self._h, _p(desc, C.c_uint8), _p(frame_offset, C.c_int32), len(frame_offset) - 1, _p(pairs, C.c_int32),
            n_pairs, max_dist, ratio, _p(off, C.c_int32), _p(oq, C.c_int32), _p(ot, C.c_int32), capacity,
            C.byref(total)))
        if out is not None:                      # the caller's arrays: views, no copy
            return off, oq[:total.value], ot[:total.value]
        return off, oq[:total.value].copy(), ot[:total.value].copy()

    # ---- resident key-frame database ----
    def kfdb_append(self, db_id: int, desc) -> int:
        """Append one key frame's descriptors (uploaded once) -> its frame number."""
        desc = _desc(desc)
        f = C.c_int(-1)
        self._check(self._lib.slamgpu_kfdb_append(self._h, db_id, _p(desc, C.c_uint8), len(desc), C.byref(f)))
        return f.value

    def kfdb_count(self, db_id: int):
        n, rows = C.c_int(0), C.c_int64(0)
        self._check(self._lib.slamgpu_kfdb_count(self._h, db_id, C.byref(n), C.byref(rows)))
        return n.value, rows.value

    def kfdb_release(self, db_id: int):
        self._check(self._lib.slamgpu_kfdb_release(self._h, db_id))

    def kfdb_match_pairs(self, db_id: int, pairs, max_dist=64.0, ratio=0.75, capacity=None, out=None):
        """match_descriptors for (query_frame, train_frame) pairs of a resident database; same return as
        match_descriptors_batch.  Only the pair list goes up."""
        pairs = np.ascontiguousarray(pairs, np.int32).reshape(-1, 2)
        n_pairs = len(pairs)
        off = np.zeros(n_pairs + 1, np.int32)
        if out is not None:
            oq, ot = out
            assert oq.dtype == np.int32 and ot.dtype == np.int32 and oq.flags.c_contiguous and ot.flags.c_contiguous
            capacity = min(len(oq), len(ot))
        else:
            if capacity is None:
                capacity = n_pairs * SLAMGPU_MAX_ROWS_DEFAULT
            oq = np.zeros(max(capacity, 1), np.int32)
            ot = np.zeros(max(capacity, 1), np.int32)
        total = C.c_int64(0)
        self._check(self._lib.slamgpu_kfdb_match_pairs(
            self._h, db_id, _p(pairs, C.c_int32), n_pairs, max_dist, ratio, _p(off, C.c_int32), _p(oq, C.c_int32),
            _p(ot, C.c_int32), capacity, C.byref(total)))
        if out is not None:                      # the caller's arrays: views, no copy
            return off, oq[:total.value], ot[:total.value]
        return off, oq[:total.value].copy(), ot[:total.value].copy()

    # ---- brute force, device buffers (raw pointers; stream-ordered, no sync) ----
    def knn2_dev(self, d_q: int, nq: int, d_t: int, nt: int, d_idx: int, d_dist: int):
        self._check(self._lib.slamgpu_knn2_dev(self._h, d_q, nq, d_t, nt, d_idx, d_dist))

    def knn2_dev_noalloc(self, d_q: int, nq: int, d_t: int, nt: int, d_idx: int, d_dist: int):
        """knn2_dev without any allocation: fixed scratch inside the library, nq and nt <= 8192 (include/slamgpu.h)."""
        self._check(self._lib.slamgpu_knn2_dev_noalloc(self._h, d_q, nq, d_t, nt, d_idx, d_dist))

    def crosscheck_dev(self, d_q: int, nq: int, d_t: int, nt: int, d_idx: int, d_dist: int):
        self._check(self._lib.slamgpu_crosscheck_dev(self._h, d_q, nq, d_t, nt, d_idx, d_dist))

    def match_descriptors_dev(self, d_q: int, nq: int, d_t: int, nt: int, max_dist: float, ratio: float,
                              d_out_train: int, d_out_count: int):
        self._check(self._lib.slamgpu_match_descriptors_dev(self._h, d_q, nq, d_t, nt, max_dist, ratio, d_out_train,
                                                            d_out_count))

    def match_descriptors_batch_dev(self, d_desc: int, d_frame_offset: int, n_frames: int, d_pairs: int, n_pairs: int,
                                    max_rows: int, max_dist: float, ratio: float, d_out_train: int, d_out_count: int):
        self._check(self._lib.slamgpu_match_descriptors_batch_dev(self._h, d_desc, d_frame_offset, n_frames, d_pairs,
                                                                  n_pairs, max_rows, max_dist, ratio, d_out_train,
                                                                  d_out_count))

    # ---- projection-guided search ----
    def frame_upload(self, frame_id: int, xy, desc, kd_perm=None, matched=None):
        xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
        desc = _desc(desc)
        assert len(desc) == len(xy), "one descriptor per keypoint"
        kd_perm = _opt(kd_perm, np.uint32)
        matched = _opt(matched, np.uint8)
        self._check(self._lib.slamgpu_frame_upload(self._h, frame_id, len(xy), _p(xy, C.c_float), _p(desc, C.c_uint8),
                                                   _p(kd_perm, C.c_uint32), _p(matched, C.c_uint8)))

    def frame_set_matched(self, frame_id: int, matched):
        matched = _opt(matched, np.uint8)
        self._check(self._lib.slamgpu_frame_set_matched(self._h, frame_id, _p(matched, C.c_uint8)))

    def frame_release(self, frame_id: int):
        self._check(self._lib.slamgpu_frame_release(self._h, frame_id))

    def frame_kd_perm(self, frame_id: int, nkp: int) -> np.ndarray:
        perm = np.zeros(nkp, np.uint32)
        self._check(self._lib.slamgpu_frame_get_kd_perm(self._h, frame_id, _p(perm, C.c_uint32)))
        return perm

    def points_upload(self, set_id: int, pos, normal, rng, obs_offset, obs_desc, obs_center=None):
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
        normal = _opt(normal, np.float32)
        rng = _opt(rng, np.float32)
        obs_offset = np.ascontiguousarray(obs_offset, np.int32)
        obs_desc = _desc(obs_desc)
        obs_center = _opt(obs_center, np.float32)
        assert len(obs_offset) == len(pos) + 1
        self._check(self._lib.slamgpu_points_upload(self._h, set_id, len(pos), _p(pos, C.c_float),
                                                    _p(normal, C.c_float), _p(rng, C.c_float),
                                                    _p(obs_offset, C.c_int32), _p(obs_desc, C.c_uint8),
                                                    _p(obs_center, C.c_float)))

    def points_append(self, set_id: int, pos, normal, rng, obs_offset, obs_desc, obs_center=None):
        """Append points at the end of a resident set; obs_offset is relative to the appended block."""
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
        normal = _opt(normal, np.float32)
        rng = _opt(rng, np.float32)
        obs_offset = np.ascontiguousarray(obs_offset, np.int32)
        obs_desc = _desc(obs_desc)
        obs_center = _opt(obs_center, np.float32)
        assert len(obs_offset) == len(pos) + 1
        self._check(self._lib.slamgpu_points_append(self._h, set_id, len(pos), _p(pos, C.c_float),
                                                    _p(normal, C.c_float), _p(rng, C.c_float),
                                                    _p(obs_offset, C.c_int32), _p(obs_desc, C.c_uint8),
                                                    _p(obs_center, C.c_float)))

    # ---- mutable map (slamgpu_points_create ...): deltas instead of re-uploads ----
    def points_create(self, set_id: int, reserve_points: int = 0, reserve_observations: int = 0,
                      reserve_observers: int = 0):
        self._check(self._lib.slamgpu_points_create(self._h, set_id, int(reserve_points), int(reserve_observations),
                                                    int(reserve_observers)))

    def observer_centers_update(self, set_id: int, observer_ids, centers):
        """New key frames, or key frames whose pose moved: Frame::camera_center() per observer id."""
        ids = np.ascontiguousarray(observer_ids, np.int32).reshape(-1)
        centers = np.ascontiguousarray(centers, np.float32).reshape(-1, 3)
        assert len(ids) == len(centers)
        self._check(self._lib.slamgpu_observer_centers_update(self._h, set_id, _p(ids, C.c_int32), len(ids),
                                                              _p(centers, C.c_float)))

    def points_append_observed(self, set_id: int, pos, obs_offset, obs_desc, obs_observer):
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
        obs_offset = np.ascontiguousarray(obs_offset, np.int32)
        obs_desc = _desc(obs_desc)
        obs_observer = np.ascontiguousarray(obs_observer, np.int32).reshape(-1)
        assert len(obs_offset) == len(pos) + 1 and len(obs_desc) == len(obs_observer)
        self._check(self._lib.slamgpu_points_append_observed(self._h, set_id, len(pos), _p(pos, C.c_float),
                                                             _p(obs_offset, C.c_int32), _p(obs_desc, C.c_uint8),
                                                             _p(obs_observer, C.c_int32)))

    def points_set_observations(self, set_id: int, idx, obs_offset, obs_desc, obs_observer):
        idx = np.ascontiguousarray(idx, np.int32).reshape(-1)
        obs_offset = np.ascontiguousarray(obs_offset, np.int32)
        obs_desc = _desc(obs_desc)
        obs_observer = np.ascontiguousarray(obs_observer, np.int32).reshape(-1)
        assert len(obs_offset) == len(idx) + 1 and len(obs_desc) == len(obs_observer)
        self._check(self._lib.slamgpu_points_set_observations(self._h, set_id, _p(idx, C.c_int32), len(idx),
                                                              _p(obs_offset, C.c_int32), _p(obs_desc, C.c_uint8),
                                                              _p(obs_observer, C.c_int32)))

    def points_set_position(self, set_id: int, idx, pos):
        idx = np.ascontiguousarray(idx, np.int32).reshape(-1)
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3)
        assert len(idx) == len(pos)
        self._check(self._lib.slamgpu_points_set_position(self._h, set_id, _p(idx, C.c_int32), len(idx),
                                                          _p(pos, C.c_float)))

    def points_remove(self, set_id: int, idx):
        idx = np.ascontiguousarray(idx, np.int32).reshape(-1)
        self._check(self._lib.slamgpu_points_remove(self._h, set_id, _p(idx, C.c_int32), len(idx)))

    def points_compact(self, set_id: int):
        """-> (remap [slots before] = new slot or -1, slots after)"""
        n_before, _ = self.points_count(set_id)
        remap = np.full(max(n_before, 1), -1, np.int32)
        n = C.c_int(0)
        self._check(self._lib.slamgpu_points_compact(self._h, set_id, _p(remap, C.c_int32), C.byref(n)))
        return remap[:n_before], n.value

    def points_live(self, set_id: int):
        """-> (live points, observation-arena entries used, of which garbage)"""
        n, used, garbage = C.c_int(0), C.c_int64(0), C.c_int64(0)
        self._check(self._lib.slamgpu_points_live(self._h, set_id, C.byref(n), C.byref(used), C.byref(garbage)))
        return n.value, used.value, garbage.value

    def points_count(self, set_id: int):
        n, o = C.c_int(0), C.c_int64(0)
        self._check(self._lib.slamgpu_points_count(self._h, set_id, C.byref(n), C.byref(o)))
        return n.value, o.value

    def points_release(self, set_id: int):
        self._check(self._lib.slamgpu_points_release(self._h, set_id))

    def points_stats(self, set_id: int, n: int):
        normal = np.zeros((n, 3), np.float32)
        rng = np.zeros((n, 2), np.float32)
        self._check(self._lib.slamgpu_points_get_stats(self._h, set_id, _p(normal, C.c_float), _p(rng, C.c_float)))
        return normal, rng

    def match_projection(self, frame_id: int, set_id: int, nkp: int, pose, params: ProjParams, skip=None, order=None,
                         debug: bool = False, n_points: int | None = None):
        """-> (out_point [nkp], out_dist [nkp]) and, with debug=True, a dict with uv / kp / d per
        visited point and the two near-boundary counters."""
        pose = np.ascontiguousarray(pose, np.float32).reshape(16)
        skip = _opt(skip, np.uint8)
        order = _opt(order, np.int32)
        n_order = 0 if order is None else len(order)
        out_point = np.full(nkp, -1, np.int32)
        out_dist = np.full(nkp, -1, np.int32)
        if not debug:
            self._check(self._lib.slamgpu_match_projection(self._h, frame_id, set_id, _p(pose, C.c_float),
                                                           C.byref(params), _p(skip, C.c_uint8), _p(order, C.c_int32),
                                                           n_order, _p(out_point, C.c_int32), _p(out_dist, C.c_int32)))
            return out_point, out_dist
        n_visit = n_order if order is not None else int(n_points)
        uv = np.zeros((n_visit, 2), np.float32)
        kp = np.zeros(n_visit, np.int32)
        d = np.zeros(n_visit, np.int32)
        stats = np.zeros(2, np.int64)
        self._check(self._lib.slamgpu_match_projection_ex(
            self._h, frame_id, set_id, _p(pose, C.c_float), C.byref(params), _p(skip, C.c_uint8),
            _p(order, C.c_int32), n_order, _p(out_point, C.c_int32), _p(out_dist, C.c_int32), _p(uv, C.c_float),
            _p(kp, C.c_int32), _p(d, C.c_int32), _p(stats, C.c_int64)))
        return out_point, out_dist, dict(uv=uv, kp=kp, d=d, near_border=int(stats[0]), near_radius=int(stats[1]))

    def match_projection_lists(self, frame_id: int, set_id: int, nkp: int, pose, params: ProjParams, matched_points,
                               order=None):
        """match_projection with frame.is_matched(point) given as an index list instead of a mask."""
        pose = np.ascontiguousarray(pose, np.float32).reshape(16)
        matched_points = np.ascontiguousarray(matched_points, np.int32).reshape(-1)
        order = _opt(order, np.int32)
        out_point = np.full(nkp, -1, np.int32)
        out_dist = np.full(nkp, -1, np.int32)
        self._check(self._lib.slamgpu_match_projection_lists(
            self._h, frame_id, set_id, _p(pose, C.c_float), C.byref(params), _p(matched_points, C.c_int32),
            len(matched_points), _p(order, C.c_int32), 0 if order is None else len(order), _p(out_point, C.c_int32),
            _p(out_dist, C.c_int32)))
        return out_point, out_dist

    def track_match(self, frame_id: int, set_id: int, nkp: int, pose, params: ProjParams, matched_points, kf_points):
        """match_key_frame -> apply -> match_map chained on the device (src/Tracker.cpp:232-248).
        -> (kf_point, kf_dist, map_point, map_dist), each [nkp]."""
        pose = np.ascontiguousarray(pose, np.float32).reshape(16)
        mp = np.ascontiguousarray(matched_points, np.int32).reshape(-1)
        kf = np.ascontiguousarray(kf_points, np.int32).reshape(-1)
        outs = [np.full(nkp, -1, np.int32) for _ in range(4)]
        self._check(self._lib.slamgpu_track_match(
            self._h, frame_id, set_id, _p(pose, C.c_float), C.byref(params), _p(mp, C.c_int32), len(mp),
            _p(kf, C.c_int32), len(kf), *[_p(o, C.c_int32) for o in outs]))
        return tuple(outs)

    # ---- bag of words ----
    def vocab_upload(self, vocab_id: int, child_offset, child_ids, desc, weight, word_id):
        """DBoW2 vocabulary tree as arrays (TemplatedVocabulary::m_nodes; node 0 = root)."""
        child_offset = np.ascontiguousarray(child_offset, np.int32)
        child_ids = np.ascontiguousarray(child_ids, np.int32)
        desc = _desc(desc)
        weight = np.ascontiguousarray(weight, np.float64)
        word_id = np.ascontiguousarray(word_id, np.int32)
        n_nodes = len(child_offset) - 1
        assert len(desc) == n_nodes and len(weight) == n_nodes and len(word_id) == n_nodes
        self._check(self._lib.slamgpu_vocab_upload(self._h, vocab_id, n_nodes, _p(child_offset, C.c_int32),
                                                   _p(child_ids, C.c_int32), _p(desc, C.c_uint8),
                                                   _p(weight, C.c_double), _p(word_id, C.c_int32)))

    def vocab_release(self, vocab_id: int):
        self._check(self._lib.slamgpu_vocab_release(self._h, vocab_id))

    def vocab_transform(self, vocab_id: int, desc):
        """-> (word [n] i32, weight [n] f64): TemplatedVocabulary::transform(feature, id, weight)."""
        desc = _desc(desc)
        word = np.full(len(desc), -1, np.int32)
        weight = np.zeros(len(desc), np.float64)
        self._check(self._lib.slamgpu_vocab_transform(self._h, vocab_id, _p(desc, C.c_uint8), len(desc),
                                                      _p(word, C.c_int32), _p(weight, C.c_double)))
        return word, weight

    def vocab_transform_dev(self, vocab_id: int, d_desc: int, n: int, d_word: int, d_weight: int = 0):
        self._check(self._lib.slamgpu_vocab_transform_dev(self._h, vocab_id, d_desc, n, d_word, d_weight or None))

    def vocab_bow(self, vocab_id: int, desc, image_offset=None, weighting=BOW_TF_IDF, norm=BOW_NORM_L1):
        """transform(features, bow) for a batch of images -> list of (ids ascending, values) per image."""
        desc = _desc(desc)
        if image_offset is None:
            image_offset = [0, len(desc)]
        off = np.ascontiguousarray(image_offset, np.int32)
        n_img = len(off) - 1
        ids = np.zeros(max(len(desc), 1), np.int32)
        vals = np.zeros(max(len(desc), 1), np.float64)
        out_off = np.zeros(n_img + 1, np.int32)
        self._check(self._lib.slamgpu_vocab_bow(self._h, vocab_id, _p(desc, C.c_uint8), _p(off, C.c_int32), n_img,
                                                weighting, norm, _p(ids, C.c_int32), _p(vals, C.c_double),
                                                _p(out_off, C.c_int32)))
        return [(ids[out_off[i]:out_off[i + 1]].copy(), vals[out_off[i]:out_off[i + 1]].copy()) for i in range(n_img)]

    def vocab_score_l1(self, query, database):
        """L1Scoring::score(query, v) for every (ids, values) vector v of `database` -> scores [n_db] f64."""
        q_ids = np.ascontiguousarray(query[0], np.int32)
        q_vals = np.ascontiguousarray(query[1], np.float64)
        off = np.zeros(len(database) + 1, np.int32)
        for i, (a, _) in enumerate(database):
            off[i + 1] = off[i] + len(a)
        db_ids = np.concatenate([np.asarray(a, np.int32) for a, _ in database]) if len(database) else np.zeros(0, np.int32)
        db_vals = np.concatenate([np.asarray(b, np.float64) for _, b in database]) if len(database) else np.zeros(0)
        db_ids, db_vals = np.ascontiguousarray(db_ids, np.int32), np.ascontiguousarray(db_vals, np.float64)
        out = np.zeros(len(database), np.float64)
        self._check(self._lib.slamgpu_vocab_score_l1(self._h, _p(q_ids, C.c_int32), _p(q_vals, C.c_double), len(q_ids),
                                                     _p(db_ids, C.c_int32), _p(db_vals, C.c_double),
                                                     _p(off, C.c_int32), len(database), _p(out, C.c_double)))
        return out

    def bowdb_append(self, db_id: int, vector):
        """Append one (ids ascending, values) vector to a resident database (created by the first append)."""
        ids = np.ascontiguousarray(vector[0], np.int32)
        vals = np.ascontiguousarray(vector[1], np.float64)
        self._check(self._lib.slamgpu_bowdb_append(self._h, db_id, _p(ids, C.c_int32), _p(vals, C.c_double), len(ids)))

    def bowdb_count(self, db_id: int) -> int:
        n = C.c_int(0)
        self._check(self._lib.slamgpu_bowdb_count(self._h, db_id, C.byref(n)))
        return n.value

    def bowdb_score_l1(self, db_id: int, query):
        q_ids = np.ascontiguousarray(query[0], np.int32)
        q_vals = np.ascontiguousarray(query[1], np.float64)
        out = np.zeros(self.bowdb_count(db_id), np.float64)
        self._check(self._lib.slamgpu_bowdb_score_l1(self._h, db_id, _p(q_ids, C.c_int32), _p(q_vals, C.c_double),
                                                     len(q_ids), _p(out, C.c_double)))
        return out

    def bowdb_release(self, db_id: int):
        self._check(self._lib.slamgpu_bowdb_release(self._h, db_id))

    # ---- PnP verification support ----
    def pnp_assemble_batch_dev(self, d_match_train: int, d_frame_offset: int, n_frames: int, d_pairs: int, n_pairs: int,
                               max_rows: int, d_row_pos: int, d_row_xy: int, d_out_object: int, d_out_image: int,
                               d_out_query: int, d_out_count: int):
        self._check(self._lib.slamgpu_pnp_assemble_batch_dev(self._h, d_match_train, d_frame_offset, n_frames, d_pairs,
                                                             n_pairs, max_rows, d_row_pos, d_row_xy, d_out_object,
                                                             d_out_image, d_out_query or None, d_out_count))

    def pnp_score_dev(self, d_object: int, d_image: int, d_corr_begin: int, d_corr_count: int, n_problems: int,
                      d_hyp_rt: int, d_hyp_problem: int, n_hyp: int, K4, reproj_error: float, d_out_count: int):
        K4 = np.ascontiguousarray(K4, np.float64).reshape(4)
        self._check(self._lib.slamgpu_pnp_score_dev(self._h, d_object, d_image, d_corr_begin, d_corr_count, n_problems,
                                                    d_hyp_rt, d_hyp_problem, n_hyp, _p(K4, C.c_double),
                                                    float(reproj_error), d_out_count))

    def pnp_mask_dev(self, d_object: int, d_image: int, d_corr_begin: int, d_corr_count: int, n_problems: int,
                     d_hyp_rt: int, d_sel_hyp: int, K4, reproj_error: float, d_out_mask: int, d_out_err: int = 0,
                     d_out_inliers: int = 0):
        K4 = np.ascontiguousarray(K4, np.float64).reshape(4)
        self._check(self._lib.slamgpu_pnp_mask_dev(self._h, d_object, d_image, d_corr_begin, d_corr_count, n_problems,
                                                   d_hyp_rt, d_sel_hyp, _p(K4, C.c_double), float(reproj_error),
                                                   d_out_mask, d_out_err or None, d_out_inliers or None))

    def pnp_ransac_score(self, objects, images, models, K4, reproj_error=4.0, confidence=0.99, model_points=5,
                         max_iters=200):
        """Score every RANSAC model of a batch of verify_pnp problems (src/LoopDetector.cpp:203-214).
        objects[p] [n_p, 3] f32, images[p] [n_p, 2] f32, models[p] [h_p, 12] f64 (R row-major, t) in iteration order
        -> list of (counts [h_p], kept model or -1, inliers, mask [n_p] u8) per problem."""
        n = len(objects)
        assert len(images) == n and len(models) == n
        co = np.zeros(n + 1, np.int32)
        ho = np.zeros(n + 1, np.int32)
        for i in range(n):
            co[i + 1] = co[i] + len(objects[i])
            ho[i + 1] = ho[i] + len(models[i])
        cat = lambda xs, w, dt: (np.ascontiguousarray(np.concatenate([np.asarray(x, dt).reshape(-1, w) for x in xs]), dt)
                                 if n else np.zeros((0, w), dt))
        obj, img, rt = cat(objects, 3, np.float32), cat(images, 2, np.float32), cat(models, 12, np.float64)
        K4 = np.ascontiguousarray(K4, np.float64).reshape(4)
        cnt = np.zeros(max(int(ho[-1]), 1), np.int32)
        best = np.zeros(max(n, 1), np.int32)
        inl = np.zeros(max(n, 1), np.int32)
        mask = np.zeros(max(int(co[-1]), 1), np.uint8)
        self._check(self._lib.slamgpu_pnp_ransac_score(self._h, _p(obj, C.c_float), _p(img, C.c_float), _p(co, C.c_int32),
                                                       n, _p(rt, C.c_double), _p(ho, C.c_int32), _p(K4, C.c_double),
                                                       float(reproj_error), float(confidence), int(model_points),
                                                       int(max_iters), _p(cnt, C.c_int32), _p(best, C.c_int32),
                                                       _p(inl, C.c_int32), _p(mask, C.c_uint8)))
        return [(cnt[ho[i]:ho[i + 1]].copy(), int(best[i]), int(inl[i]), mask[co[i]:co[i + 1]].copy()) for i in range(n)]

    def match_projection_batch(self, frame_ids, nkps, set_id: int, poses, params: ProjParams, skip=None, order=None):
        """Frozen-snapshot batch over key frames -> list of (out_point, out_dist) per frame."""
        frame_ids = [int(f) for f in frame_ids]
        ids = (C.c_int * len(frame_ids))(*frame_ids)
        poses = np.ascontiguousarray(poses, np.float32).reshape(len(frame_ids), 16)
        skip = _opt(skip, np.uint8)
        order = _opt(order, np.int32)
        n_order = 0 if order is None else len(order)
        total = int(sum(nkps))
        out_point = np.full(max(total, 1), -1, np.int32)
        out_dist = np.full(max(total, 1), -1, np.int32)
        self._check(self._lib.slamgpu_match_projection_batch(
            self._h, ids, len(frame_ids), set_id, _p(poses, C.c_float), C.byref(params), _p(skip, C.c_uint8),
            _p(order, C.c_int32), n_order, _p(out_point, C.c_int32), _p(out_dist, C.c_int32)))
        res, o = [], 0
        for n in nkps:
            res.append((out_point[o:o + n].copy(), out_dist[o:o + n].copy()))
            o += n
        return res
