"""Host-side mirror of the reference's matcher interface, on top of libslamgpu.

Same names, argument meaning and empty-result behaviour as the reference (paths relative to the
reference tree):

  MapMatcher                      src/MapMatcher.h:16-38
    .match_map(frame, map)                        :21   -> [MapPointMatch]
    .match_key_frame(frame, map, key_frame)       :24
    .match_for_fuse(frame, points)                :27
    .match_descriptors(frame, key_frame)          :30
  match_features(prev, cur, max_distance)         src/features/FeatureExtractor.cpp:15-30
  ExtractedFeatures / FeatureMatch / MapPointMatch  src/features/FeatureExtractor.h:15-38
  Camera / Frame / KeyFrame / MapPoint / Map      read accessors only

The data-model classes below hold exactly what the matcher reads; they exist so that tests read
like the reference's call sites.  All matching arithmetic runs in the CUDA library -- this module
only flattens objects to POD arrays (what the C++ adapter in INTEGRATION.md does) and maps
indices back to objects.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import binding as L

NORM_HAMMING = 6          # cv::NORM_HAMMING
ORB_MAX_DISTANCE = 64.0   # src/features/OrbFeatureExtractor.h:14-17

# src/MapMatcher.cpp:12-18
SEARCH_RADIUS = 20.0
MIN_VIEWING_ANGLE_COSINE = 0.5
MAX_NEARER_RATIO = 2.0
MAX_FURTHER_RATIO = 1.25
MATCH_RATIO = 0.75


@dataclass
class ExtractedFeatures:
    keypoints: np.ndarray      # [n,2] f32 (cv::KeyPoint::pt)
    descriptors: np.ndarray    # [n,32] u8


@dataclass
class FeatureMatch:
    train_index: int
    query_index: int


@dataclass
class Camera:
    fx: float
    fy: float
    cx: float
    cy: float
    width: int
    height: int

    def get_width(self):
        return self.width

    def get_height(self):
        return self.height


class MapPoint:
    """position + observations (KeyFrame -> keypoint index, in insertion = iteration order)."""

    def __init__(self, position):
        self._position = np.asarray(position, np.float32).reshape(3)
        self._observations: dict = {}

    def position(self):
        return self._position

    def add_observation(self, key_frame, index: int):
        self._observations[key_frame] = int(index)

    def observations(self):
        return self._observations

    def is_observed_by(self, key_frame) -> bool:
        return key_frame in self._observations


@dataclass
class MapPointMatch:
    point: MapPoint
    keypoint_index: int


class Frame:
    def __init__(self, index: int, features: ExtractedFeatures, pose=None):
        self.index = index
        self._features = ExtractedFeatures(np.ascontiguousarray(features.keypoints, np.float32).reshape(-1, 2),
                                           np.ascontiguousarray(features.descriptors, np.uint8).reshape(-1, 32))
        self._pose = np.eye(4, dtype=np.float32) if pose is None else np.asarray(pose, np.float32).reshape(4, 4)
        self._map_matches: list = [None] * len(self._features.keypoints)
        self._matched_points: set = set()

    def features(self):
        return self._features

    def pose(self):
        return self._pose

    def set_pose(self, pose):
        self._pose = np.asarray(pose, np.float32).reshape(4, 4)

    def camera_center(self):
        """src/Frame.cpp:39-42 in float32 (a0 + (a1 + a2) per component, like the library)."""
        R, t = self._pose[:3, :3], self._pose[:3, 3]
        c = np.zeros(3, np.float32)
        for i in range(3):
            p = [np.float32(-R[k, i]) * np.float32(t[k]) for k in range(3)]
            c[i] = np.float32(p[0] + np.float32(p[1] + p[2]))
        return c

    def descriptor(self, index: int):
        return self._features.descriptors[index]

    def add_map_match(self, match: MapPointMatch):
        self._map_matches[match.keypoint_index] = match.point
        self._matched_points.add(id(match.point))

    def is_matched(self, what) -> bool:
        if isinstance(what, MapPoint):
            return id(what) in self._matched_points
        return self._map_matches[int(what)] is not None

    def map_matches(self):
        """Ascending keypoint order, like Frame::MapPointIterator (src/Frame.cpp:154-174)."""
        return [MapPointMatch(p, i) for i, p in enumerate(self._map_matches) if p is not None]

    def num_map_matches(self):
        return sum(p is not None for p in self._map_matches)


class KeyFrame(Frame):
    def __hash__(self):
        return id(self)

    def __eq__(self, other):
        return self is other


class Map:
    """Owns points; iteration order = insertion order (src/Map.cpp:164-190)."""

    def __init__(self):
        self._points: list = []

    def add_point(self, point: MapPoint):
        self._points.append(point)
        return point

    def __iter__(self):
        return iter(self._points)

    def __len__(self):
        return len(self._points)


def match_features(ctx: L.Context, prev_features: ExtractedFeatures, features: ExtractedFeatures,
                   max_distance: float = ORB_MAX_DISTANCE):
    """BaseFeatureExtractor::match_features (src/features/FeatureExtractor.cpp:15-30)."""
    tr, qu = ctx.match_features(prev_features.descriptors, features.descriptors, max_distance)
    return [FeatureMatch(int(t), int(q)) for t, q in zip(tr, qu)]


@dataclass
class _Flat:
    points: list
    pos: np.ndarray
    obs_offset: np.ndarray
    obs_desc: np.ndarray
    obs_center: np.ndarray
    order: np.ndarray = field(default=None)


def _flatten_points(points) -> _Flat:
    """What the adapter uploads per map state: position, one descriptor and one observer camera
    centre per observation, in the host's observations() iteration order (src/MapMatcher.cpp:86)."""
    uniq, seen, order = [], {}, []
    for p in points:
        if p is None:
            order.append(-1)                      # match_for_fuse nullptr, :119-121
            continue
        k = seen.get(id(p))
        if k is None:
            k = seen[id(p)] = len(uniq)
            uniq.append(p)
        order.append(k)
    n = len(uniq)
    pos = np.zeros((n, 3), np.float32)
    off = np.zeros(n + 1, np.int32)
    desc, cent = [], []
    centers = {}
    for i, p in enumerate(uniq):
        pos[i] = p.position()
        for kf, idx in p.observations().items():
            desc.append(kf.descriptor(idx))
            c = centers.get(id(kf))
            if c is None:
                c = centers[id(kf)] = kf.camera_center()
            cent.append(c)
        off[i + 1] = len(desc)
    obs_desc = np.asarray(desc, np.uint8).reshape(-1, 32)
    obs_center = np.asarray(cent, np.float32).reshape(-1, 3)
    return _Flat(uniq, pos, off, obs_desc, obs_center, np.asarray(order, np.int32))


class MapMatcher:
    """Drop-in for slam::MapMatcher: same four entry points, results as [MapPointMatch]."""

    _FRAME_SLOT = 0
    _SET_SLOT = 0

    def __init__(self, camera: Camera, max_descriptor_distance: float = ORB_MAX_DISTANCE, norm_type=NORM_HAMMING,
                 ctx: L.Context | None = None, search_radius: float = SEARCH_RADIUS, match_ratio: float = MATCH_RATIO):
        if norm_type != NORM_HAMMING:
            raise ValueError("libslamgpu implements NORM_HAMMING (ORB) matching only")
        self.camera = camera
        self.max_descriptor_distance = float(max_descriptor_distance)
        self.ctx = ctx if ctx is not None else L.Context(0)
        self.search_radius = float(search_radius)
        self.match_ratio = float(match_ratio)

    def _params(self, replace: bool) -> L.ProjParams:
        c = self.camera
        return L.ProjParams(c.fx, c.fy, c.cx, c.cy, c.width, c.height, self.search_radius,
                            self.max_descriptor_distance, MIN_VIEWING_ANGLE_COSINE, MAX_NEARER_RATIO,
                            MAX_FURTHER_RATIO, 1 if replace else 0)

    def _project(self, frame: Frame, points, skip_fn, replace: bool):
        flat = _flatten_points(points)
        feats = frame.features()
        nkp = len(feats.keypoints)
        if nkp == 0:
            return []
        matched = np.array([frame.is_matched(i) for i in range(nkp)], np.uint8)
        skip = np.array([1 if skip_fn(p) else 0 for p in flat.points], np.uint8)
        self.ctx.frame_upload(self._FRAME_SLOT, feats.keypoints, feats.descriptors, None, matched)
        self.ctx.points_upload(self._SET_SLOT, flat.pos, None, None, flat.obs_offset, flat.obs_desc, flat.obs_center)
        pose = np.ascontiguousarray(frame.pose().flatten(order="F"))     # Eigen::Matrix4f is column-major
        out_point, _ = self.ctx.match_projection(self._FRAME_SLOT, self._SET_SLOT, nkp, pose, self._params(replace),
                                                 skip, flat.order)
        # accepted_matches: ascending keypoint index (src/MapMatcher.cpp:34-43)
        return [MapPointMatch(flat.points[int(p)], kp) for kp, p in enumerate(out_point) if p >= 0]

    def match_map(self, frame: Frame, map_: Map):
        return self._project(frame, list(map_), lambda p: frame.is_matched(p), False)

    def match_key_frame(self, frame: Frame, map_: Map, key_frame: KeyFrame):
        return self._project(frame, list(map_),
                             lambda p: frame.is_matched(p) or not p.is_observed_by(key_frame), False)

    def track(self, frame: Frame, map_: Map, last_key_frame: KeyFrame):
        """Tracker::match_with_last_key_frame followed by Tracker::match_with_map
        (src/Tracker.cpp:232-248) as one chained device call (slamgpu_track_match): both match lists
        are applied to `frame` with add_map_match, exactly as the two reference loops do.
        -> (key-frame matches, map matches)."""
        points = list(map_)
        flat = _flatten_points(points)
        feats = frame.features()
        nkp = len(feats.keypoints)
        if nkp == 0:
            return [], []
        matched = np.array([frame.is_matched(i) for i in range(nkp)], np.uint8)
        self.ctx.frame_upload(self._FRAME_SLOT, feats.keypoints, feats.descriptors, None, matched)
        self.ctx.points_upload(self._SET_SLOT, flat.pos, None, None, flat.obs_offset, flat.obs_desc, flat.obs_center)
        matched_points = [i for i, p in enumerate(flat.points) if frame.is_matched(p)]
        kf_points = [i for i, p in enumerate(flat.points) if p.is_observed_by(last_key_frame)]   # :169, map order
        pose = np.ascontiguousarray(frame.pose().flatten(order="F"))
        kf_pt, _, map_pt, _ = self.ctx.track_match(self._FRAME_SLOT, self._SET_SLOT, nkp, pose, self._params(False),
                                                   matched_points, kf_points)
        kf_matches = [MapPointMatch(flat.points[int(p)], kp) for kp, p in enumerate(kf_pt) if p >= 0]
        map_matches = [MapPointMatch(flat.points[int(p)], kp) for kp, p in enumerate(map_pt) if p >= 0]
        for m in kf_matches:
            frame.add_map_match(m)
        for m in map_matches:
            frame.add_map_match(m)
        return kf_matches, map_matches

    def match_for_fuse(self, frame: Frame, points):
        return self._project(frame, list(points), lambda p: frame.is_matched(p), True)

    def match_descriptors(self, frame: Frame, key_frame: KeyFrame):
        matches = key_frame.map_matches()                              # :135-138
        q = frame.features().descriptors
        if not matches or len(q) == 0:
            return []                                                  # :139-141
        train = np.stack([key_frame.descriptor(m.keypoint_index) for m in matches])
        qi, ti = self.ctx.match_descriptors(q, train, self.max_descriptor_distance, self.match_ratio)
        return [MapPointMatch(matches[int(t)].point, int(qq)) for qq, t in zip(qi, ti)]


class OrbVocabulary:
    """Host-side mirror of what LoopDetector uses of DBoW2's OrbVocabulary (src/LoopDetector.cpp:26,85-92,369):
    transform(features) -> bag-of-words vector and score(a, b), on a vocabulary tree resident on the device.
    A vector is (ids ascending, values) -- the contents of DBoW2::BowVector's std::map in order."""

    _next_id = 100

    def __init__(self, ctx: L.Context, child_offset, child_ids, desc, weight, word_id,
                 weighting: int = L.BOW_TF_IDF, norm: int = L.BOW_NORM_L1):
        self.ctx, self.weighting, self.norm = ctx, weighting, norm
        self.vocab_id = OrbVocabulary._next_id
        OrbVocabulary._next_id += 1
        ctx.vocab_upload(self.vocab_id, child_offset, child_ids, desc, weight, word_id)

    def transform(self, descriptors):
        """bow_of (src/LoopDetector.cpp:85-92): empty features give an empty vector."""
        d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        if len(d) == 0:
            return np.zeros(0, np.int32), np.zeros(0, np.float64)
        return self.ctx.vocab_bow(self.vocab_id, d, None, self.weighting, self.norm)[0]

    def score(self, a, b) -> float:
        return float(self.ctx.vocab_score_l1(a, [b])[0])

    def close(self):
        self.ctx.vocab_release(self.vocab_id)


class KeyFrameDatabase:
    """The key frames LoopDetector keeps (src/LoopDetector.cpp:314-330) resident on the device: add() uploads a key
    frame's descriptors once; match_pairs() is match_descriptors (src/MapMatcher.cpp:129-163) for a list of
    (query key frame, candidate key frame) numbers -- verify_pnp's first step for every retrieved candidate."""

    _next_id = 100

    def __init__(self, ctx: L.Context):
        self.ctx = ctx
        self.db_id = KeyFrameDatabase._next_id
        KeyFrameDatabase._next_id += 1
        self._n = 0
        self._max_rows = 0

    def add(self, descriptors) -> int:
        f = self.ctx.kfdb_append(self.db_id, descriptors)
        self._n += 1
        self._max_rows = max(self._max_rows, len(descriptors))
        return f

    def __len__(self):
        return self._n

    def match_pairs(self, pairs, max_dist=64.0, ratio=0.75, out=None):
        pairs = np.asarray(pairs, np.int32).reshape(-1, 2)
        return self.ctx.kfdb_match_pairs(self.db_id, pairs, max_dist, ratio,
                                         capacity=len(pairs) * max(self._max_rows, 1), out=out)

    def close(self):
        if self._n:
            self.ctx.kfdb_release(self.db_id)
            self._n = 0


class BowDatabase:
    """LoopDetector::Impl::bows kept on the device: add() per key frame, scores() of a query against all."""

    _next_id = 100

    def __init__(self, ctx: L.Context):
        self.ctx = ctx
        self.db_id = BowDatabase._next_id
        BowDatabase._next_id += 1
        self._n = 0

    def add(self, vector) -> int:
        self.ctx.bowdb_append(self.db_id, vector)
        self._n += 1
        return self._n - 1

    def scores(self, query) -> np.ndarray:
        """score_candidates' inner loop (src/LoopDetector.cpp:369) for every stored key frame."""
        if self._n == 0:
            return np.zeros(0, np.float64)
        return self.ctx.bowdb_score_l1(self.db_id, query)

    def close(self):
        if self._n:
            self.ctx.bowdb_release(self.db_id)
            self._n = 0
