"""The C++ adapter (racing-slam_b200/adapter/MapMatcherGpu.{h,cpp}) compiled against the mock of the
reference's data model (tests/adapter_mock/) and run end to end: world -> MapMatcherGpu -> libslamgpu ->
[MapPointMatch], compared with the oracle fed the same world in the host's observation order."""
import os
import struct
import subprocess

import numpy as np
import pytest

from conftest import ROOT

MOCK = os.path.join(ROOT, "tests", "adapter_mock")
DRIVER = os.path.join(MOCK, "adapter_driver")


def build_driver():
    r = subprocess.run(["make", "-C", MOCK], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert os.path.exists(DRIVER)


def make_world(pkg, seed=71, n_points=1200, nkp=500, n_kf=4, kf_kp=300):
    syn = pkg.synthetic
    rng = np.random.default_rng(seed)
    cam = syn.KITTI
    sc = syn.projection_scene(n_points, nkp, seed)
    w = dict(cam=cam, kfs=[], points=[], kf_matches=[], frame=None, frame_matches=[], fuse=[], required=1)
    for k in range(n_kf):
        T = syn.make_pose(rng, 3.0, 0.8)
        w["kfs"].append(dict(xy=syn.keypoints_integer(kf_kp, cam["width"], cam["height"], seed + k),
                             desc=rng.integers(0, 256, (kf_kp, 32), dtype=np.uint8), pose=syn.pose_colmajor(T)))
    for i in range(n_points):
        obs, used = [], set()
        for o in range(sc.obs_offset[i], sc.obs_offset[i + 1]):
            k = int(rng.integers(0, n_kf))
            if k in used:                       # observations() is keyed by key frame: one per key frame
                continue
            used.add(k)
            idx = int(rng.integers(0, kf_kp))
            w["kfs"][k]["desc"][idx] = sc.obs_desc[o]
            obs.append((k, idx))
            if rng.random() < 0.3:
                w["kf_matches"].append((k, idx, i))
        w["points"].append(dict(pos=sc.pos[i], obs=obs))
    # one map point per key-frame keypoint
    seen, km = set(), []
    for k, idx, i in w["kf_matches"]:
        if (k, idx) not in seen:
            seen.add((k, idx))
            km.append((k, idx, i))
    w["kf_matches"] = km
    w["frame"] = dict(xy=sc.kp_xy, desc=sc.kp_desc, pose=sc.pose)
    used_p = set()
    for kp in np.nonzero(sc.kp_matched)[0][:25]:
        pi = int(rng.integers(0, n_points))
        if pi not in used_p:
            used_p.add(pi)
            w["frame_matches"].append((int(kp), pi))
    fuse = list(range(0, n_points, 3)) + [-1, -1] + list(range(40))
    w["fuse"] = fuse
    return w


def write_world(w, path):
    with open(path, "wb") as f:
        c = w["cam"]
        f.write(struct.pack("<ffffii", c["fx"], c["fy"], c["cx"], c["cy"], c["width"], c["height"]))
        f.write(struct.pack("<i", len(w["kfs"])))
        for kf in w["kfs"]:
            f.write(struct.pack("<i", len(kf["xy"])))
            f.write(np.ascontiguousarray(kf["xy"], np.float32).tobytes())
            f.write(np.ascontiguousarray(kf["desc"], np.uint8).tobytes())
            f.write(np.ascontiguousarray(kf["pose"], np.float32).tobytes())
        f.write(struct.pack("<i", len(w["points"])))
        for p in w["points"]:
            f.write(np.ascontiguousarray(p["pos"], np.float32).tobytes())
            f.write(struct.pack("<i", len(p["obs"])))
            for k, idx in p["obs"]:
                f.write(struct.pack("<ii", k, idx))
        f.write(struct.pack("<i", len(w["kf_matches"])))
        for k, idx, i in w["kf_matches"]:
            f.write(struct.pack("<iii", k, idx, i))
        fr = w["frame"]
        f.write(struct.pack("<i", len(fr["xy"])))
        f.write(np.ascontiguousarray(fr["xy"], np.float32).tobytes())
        f.write(np.ascontiguousarray(fr["desc"], np.uint8).tobytes())
        f.write(np.ascontiguousarray(fr["pose"], np.float32).tobytes())
        f.write(struct.pack("<i", len(w["frame_matches"])))
        for kp, pi in w["frame_matches"]:
            f.write(struct.pack("<ii", kp, pi))
        f.write(struct.pack("<i", len(w["fuse"])))
        f.write(np.asarray(w["fuse"], np.int32).tobytes())
        f.write(struct.pack("<i", w["required"]))


def parse_output(path):
    lines = open(path).read().split("\n")
    res, i = {}, 0
    while i < len(lines) and lines[i].strip():
        name, n = lines[i].split()
        n = int(n)
        if name == "frame_matches_after_track":
            res[name] = n
            i += 1
            continue
        rows = [tuple(int(x) for x in lines[i + 1 + r].split()) for r in range(n)]
        res[name] = rows
        i += 1 + n
    return res


def test_adapter_builds_against_mock_and_fails_loudly_without_gpu(pkg, tmp_path):
    build_driver()
    if pkg.binding.device_count() > 0:
        pytest.skip("a CUDA device is present; the run is covered by the gpu test")
    w = make_world(pkg, n_points=50, nkp=30, n_kf=2, kf_kp=40)
    write_world(w, tmp_path / "world.bin")
    r = subprocess.run([DRIVER, str(tmp_path / "world.bin"), str(tmp_path / "out.txt")], capture_output=True, text=True)
    assert r.returncode == 1 and "libslamgpu" in r.stderr       # no CPU fallback: the adapter throws


@pytest.mark.gpu
def test_adapter_end_to_end_vs_oracle(pkg, oracle, tmp_path):
    build_driver()
    w = make_world(pkg)
    write_world(w, tmp_path / "world.bin")
    r = subprocess.run([DRIVER, str(tmp_path / "world.bin"), str(tmp_path / "out.txt")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    got = parse_output(tmp_path / "out.txt")

    # oracle inputs in the host's observation order (what the C++ unordered_map iterated)
    kfs, cam, fr = w["kfs"], w["cam"], w["frame"]
    centers = [oracle.camera_center(kf["pose"]) for kf in kfs]
    n = len(w["points"])
    pos = np.stack([p["pos"] for p in w["points"]]).astype(np.float32)
    off = np.zeros(n + 1, np.int32)
    od, oc = [], []
    observers = []
    for i, row in enumerate(got["obs_order"]):
        cnt, rest = row[0], row[1:]
        obs = [(rest[2 * j], rest[2 * j + 1]) for j in range(cnt)]
        assert sorted(obs) == sorted(w["points"][i]["obs"])
        observers.append({k for k, _ in obs})
        for k, idx in obs:
            od.append(kfs[k]["desc"][idx])
            oc.append(centers[k])
        off[i + 1] = len(od)
    od = np.asarray(od, np.uint8).reshape(-1, 32)
    oc = np.asarray(oc, np.float32).reshape(-1, 3)
    normal, rng_ = oracle.point_stats(pos, off, oc)
    nkp = len(fr["xy"])
    matched = np.zeros(nkp, np.uint8)
    matched_points = set()
    for kp, pi in w["frame_matches"]:
        matched[kp] = 1
        matched_points.add(pi)
    perm = oracle.kd_build(fr["xy"])

    def run(skip, order, replace):
        prm = oracle.ProjParams(cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["width"], cam["height"], 20.0, 64.0,
                                0.5, 2.0, 1.25, replace)
        res = oracle.match_projection(fr["xy"], fr["desc"], perm, matched, pos, normal, rng_, off, od, skip, order,
                                      fr["pose"], prm, want_stats=False)
        return [(int(p), kp) for kp, p in enumerate(res.point) if p >= 0]

    skip_m = np.array([1 if i in matched_points else 0 for i in range(n)], np.uint8)
    assert got["match_map"] == run(skip_m, None, 0) and len(got["match_map"]) > 20
    req = w["required"]
    skip_k = np.array([1 if (i in matched_points or req not in observers[i]) else 0 for i in range(n)], np.uint8)
    assert got["match_key_frame"] == run(skip_k, None, 0) and len(got["match_key_frame"]) > 3
    assert got["match_for_fuse"] == run(skip_m, np.asarray(w["fuse"], np.int32), 1) and len(got["match_for_fuse"]) > 10

    kfm = sorted((idx, i) for k, idx, i in w["kf_matches"] if k == req)      # ascending keypoint index
    train = np.stack([kfs[req]["desc"][idx] for idx, _ in kfm])
    oq, ot = oracle.match_descriptors(fr["desc"], train, 64.0, 0.75)
    assert got["match_descriptors"] == [(kfm[int(t)][1], int(q)) for q, t in zip(oq, ot)]
    # match_descriptors_batch: every key frame as a candidate, the same list as one match_descriptors each
    assert got[f"mdb{req}"] == got["match_descriptors"]
    n_lists = 0
    for k in range(len(kfs)):
        kfm_k = sorted((idx, i) for kk, idx, i in w["kf_matches"] if kk == k)
        if not kfm_k:
            assert got[f"mdb{k}"] == []
            continue
        train_k = np.stack([kfs[k]["desc"][idx] for idx, _ in kfm_k])
        bq, bt = oracle.match_descriptors(fr["desc"], train_k, 64.0, 0.75)
        assert got[f"mdb{k}"] == [(kfm_k[int(t)][1], int(q)) for q, t in zip(bq, bt)], k
        n_lists += 1
    assert n_lists >= 2
    o_t, o_q = oracle.match_features(kfs[0]["desc"], fr["desc"], 64.0)
    assert got["match_features"] == list(zip(o_t.tolist(), o_q.tolist()))

    # MapMatcherGpu::track = match_key_frame, add_map_match loop, match_map, add_map_match loop
    # (src/Tracker.cpp:232-248): the first stage equals the stand-alone call, the second sees its matches
    assert got["track_key_frame"] == got["match_key_frame"]
    skip_after = skip_m.copy()
    for pi, kp in got["track_key_frame"]:
        matched[kp] = 1
        skip_after[pi] = 1
    assert got["track_map"] == run(skip_after, None, 0) and len(got["track_map"]) > 10
    assert got["frame_matches_after_track"] == len(w["frame_matches"]) + len(got["track_key_frame"]) + len(got["track_map"])
