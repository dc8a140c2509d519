"""Independent numpy integer model of cv::BFMatcher's kNN / cross-check semantics.

TEST INFRASTRUCTURE ONLY (see oracle/slam_oracle.cpp header).  A second, independent statement
of the brute-force rules, used to cross-check both the C++ oracle and the cv2 golden vectors:

  * kNN      = first k of a STABLE sort by (distance, trainIdx)
  * crosscheck keeps (q,t) iff t = argmin_t d(q,.) and q = argmin_q d(.,t), lowest index winning
    ties in both directions; failed rows come back empty.

Reference call sites: src/MapMatcher.cpp:147-148, src/features/FeatureExtractor.cpp:20-21.
"""
from __future__ import annotations

import numpy as np

_POP8 = np.array([bin(i).count("1") for i in range(256)], dtype=np.int32)


def hamming_matrix(q: np.ndarray, t: np.ndarray, chunk: int = 512) -> np.ndarray:
    """[nq, nt] int32 Hamming distances between 32-byte rows (cv::norm NORM_HAMMING)."""
    q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32)
    t = np.ascontiguousarray(t, np.uint8).reshape(-1, 32)
    out = np.zeros((len(q), len(t)), np.int32)
    for s in range(0, len(q), chunk):
        x = q[s:s + chunk, None, :] ^ t[None, :, :]
        out[s:s + chunk] = _POP8[x].sum(axis=2)
    return out


def knn2(q, t):
    d = hamming_matrix(q, t)
    nq, nt = d.shape
    idx = np.full((nq, 2), -1, np.int32)
    dist = np.full((nq, 2), -1, np.int32)
    if nt == 0:
        return idx, dist
    order = np.argsort(d, axis=1, kind="stable")[:, :2]
    k = order.shape[1]
    idx[:, :k] = order
    dist[:, :k] = np.take_along_axis(d, order, axis=1)
    return idx, dist


def crosscheck(q, t):
    d = hamming_matrix(q, t)
    nq, nt = d.shape
    idx = np.full(nq, -1, np.int32)
    dist = np.full(nq, -1, np.int32)
    if nt == 0 or nq == 0:
        return idx, dist
    row = d.argmin(axis=1)            # numpy argmin returns the FIRST minimum = lowest index
    col = d.argmin(axis=0)
    ok = col[row] == np.arange(nq)
    idx[ok] = row[ok]
    dist[ok] = d[np.arange(nq), row][ok]
    return idx, dist


def match_descriptors(q, t, max_dist=64.0, ratio=0.75):
    """src/MapMatcher.cpp:145-161 on kNN rows -> (query_idx, train_idx)."""
    q = np.asarray(q).reshape(-1, 32)
    t = np.asarray(t).reshape(-1, 32)
    if len(q) == 0 or len(t) == 0:
        return np.zeros(0, np.int32), np.zeros(0, np.int32)
    idx, dist = knn2(q, t)
    d0 = dist[:, 0].astype(np.float32)
    keep = ~(d0 > np.float32(max_dist))
    if len(t) >= 2:
        d1 = dist[:, 1].astype(np.float32)
        keep &= ~(d0 > np.float32(ratio) * d1)
    qi = np.nonzero(keep)[0].astype(np.int32)
    return qi, idx[qi, 0]


def match_features(prev, cur, max_dist=64.0):
    """src/features/FeatureExtractor.cpp:15-30 -> (train_idx, query_idx)."""
    idx, dist = crosscheck(cur, prev)
    keep = (idx >= 0) & (dist.astype(np.float32) < np.float32(max_dist))
    qi = np.nonzero(keep)[0].astype(np.int32)
    return idx[qi], qi
