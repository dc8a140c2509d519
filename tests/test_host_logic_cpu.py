"""CPU tests of the host logic: multi-GPU sharding (incl. a world_size-2 gloo run), the synthetic
generators and the flattening the matcher mirror performs."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT


def test_all_pairs_is_lower_triangle(pkg):
    p = pkg.sharding.all_pairs(6)
    assert len(p) == 15 and (p[:, 0] > p[:, 1]).all()
    assert len({tuple(x) for x in p.tolist()}) == 15


@pytest.mark.parametrize("world", [1, 2, 4, 8])
def test_shard_pairs_partitions_exactly(pkg, world):
    pairs = pkg.sharding.all_pairs(40)
    parts = [pkg.sharding.shard_pairs(pairs, r, world, block=16) for r in range(world)]
    allp = np.concatenate(parts)
    assert len(allp) == len(pairs)
    assert {tuple(x) for x in allp.tolist()} == {tuple(x) for x in pairs.tolist()}
    sizes = [len(p) for p in parts]
    assert max(sizes) - min(sizes) <= 16


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_shard_keyframes_round_robin(pkg, world):
    got = np.sort(np.concatenate([pkg.sharding.shard_keyframes(20, r, world) for r in range(world)]))
    assert np.array_equal(got, np.arange(20))


def _gloo_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import racing_slam_b200 as pkg
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        pairs = pkg.sharding.all_pairs(12)
        idx = np.nonzero((np.arange(len(pairs)) // 4) % world == rank)[0]
        mine = pkg.sharding.shard_pairs(pairs, rank, world, block=4)
        assert np.array_equal(mine, pairs[idx])
        # stand-in per-pair result (the GPU would produce match counts): a function of the pair
        local = mine[:, 0] * 100 + mine[:, 1]
        full = pkg.sharding.gather_counts(local, idx, len(pairs))
        # the one exchange of the batched workloads: the database from rank 0 to everyone
        db = pkg.synthetic.keyframe_db(3, 16, 5) if rank == 0 else None
        got, _ = pkg.sharding.replicate(db, src=0)
        assert got.shape == (3, 16, 32) and got.dtype == np.uint8
        assert np.array_equal(got, pkg.synthetic.keyframe_db(3, 16, 5))
        empty, _ = pkg.sharding.replicate(np.zeros((0, 32), np.uint8) if rank == 0 else None, src=0)
        assert empty.shape == (0, 32)
        q.put((rank, full.tolist()))
    finally:
        dist.destroy_process_group()


def test_gloo_world2_shard_and_gather():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    import racing_slam_b200 as pkg
    pairs = pkg.sharding.all_pairs(12)
    want = (pairs[:, 0] * 100 + pairs[:, 1]).tolist()
    assert res[0] == want and res[1] == want


def test_synthetic_shapes(pkg):
    syn = pkg.synthetic
    q, t = syn.bf_correlated(100, 80, 1)
    assert q.shape == (100, 32) and t.shape == (80, 32) and q.dtype == np.uint8
    sc = syn.projection_scene(500, 200, 2)
    assert sc.pos.shape == (500, 3) and sc.obs_offset[-1] == len(sc.obs_desc) == len(sc.obs_center)
    assert sc.pose.shape == (16,) and sc.kp_xy.shape == (200, 2)
    db = syn.keyframe_db(5, 64, 3)
    assert db.shape == (5, 64, 32)
    f = syn.flip_bits(q, 7, np.random.default_rng(0))
    assert (np.unpackbits(f ^ q, axis=1).sum(1) == 7).all()


def test_matcher_flattening_keeps_host_observation_order(pkg):
    m = pkg.matcher
    rng = np.random.default_rng(0)
    kfs = [m.KeyFrame(i, m.ExtractedFeatures(rng.uniform(0, 50, (10, 2)), rng.integers(0, 256, (10, 32), dtype=np.uint8)))
           for i in range(3)]
    p0, p1 = m.MapPoint([1, 2, 3]), m.MapPoint([4, 5, 6])
    p0.add_observation(kfs[2], 4)
    p0.add_observation(kfs[0], 1)
    p1.add_observation(kfs[1], 9)
    flat = m._flatten_points([p1, None, p0, p1])
    assert flat.order.tolist() == [0, -1, 1, 0]
    assert flat.obs_offset.tolist() == [0, 1, 3]
    assert np.array_equal(flat.obs_desc[0], kfs[1].descriptor(9))
    assert np.array_equal(flat.obs_desc[1], kfs[2].descriptor(4))   # insertion order, not key order
    assert np.array_equal(flat.obs_desc[2], kfs[0].descriptor(1))
    f = m.Frame(7, m.ExtractedFeatures(np.zeros((3, 2)), np.zeros((3, 32), np.uint8)))
    f.add_map_match(m.MapPointMatch(p0, 2))
    assert f.is_matched(p0) and not f.is_matched(p1) and f.is_matched(2) and not f.is_matched(0)
    assert [mm.keypoint_index for mm in f.map_matches()] == [2]


def test_host_restatements_equal_the_oracle_bit_for_bit(pkg, oracle):
    """slamgpu_host_project / _camera_center / _point_stats (what MapMatcherGpu::self_check / audit and
    tools/eigen_order_check.cpp compare with the reference's Eigen code) are the oracle's operation order."""
    syn = pkg.synthetic
    for cam, seed in ((syn.KITTI, 5), (syn.HD1080, 6)):     # HD1080: u >= 1024 px, one ulp = 1.2e-4 px
        sc = syn.projection_scene(3000, 100, seed, cam=cam)
        prm = pkg.ProjParams(cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["width"], cam["height"], 20.0, 64.0, 0.5,
                             2.0, 1.25, 0)
        uv = pkg.binding.host_project(sc.pose, prm, sc.pos)
        ref = oracle.project(sc.pose, cam["fx"], cam["fy"], cam["cx"], cam["cy"], sc.pos)
        assert np.array_equal(uv.view(np.uint32), ref.view(np.uint32))
        assert (uv[:, 0] >= 1024).sum() > 100 or cam is syn.KITTI
        c = pkg.binding.host_camera_center(sc.pose)
        assert np.array_equal(c.view(np.uint32), oracle.camera_center(sc.pose).view(np.uint32))
        normal, rng_ = oracle.point_stats(sc.pos, sc.obs_offset, sc.obs_center)
        for i in range(0, 3000, 7):
            n, r = pkg.binding.host_point_stats(sc.pos[i], sc.obs_center[sc.obs_offset[i]:sc.obs_offset[i + 1]])
            assert np.array_equal(n.view(np.uint32), normal[i].view(np.uint32))
            assert np.array_equal(r.view(np.uint32), rng_[i].view(np.uint32))
    n, r = pkg.binding.host_point_stats(np.zeros(3, np.float32), np.zeros((0, 3), np.float32))   # no observation
    assert np.array_equal(n, np.zeros(3, np.float32)) and r[0] == np.finfo(np.float32).max and r[1] == 0.0
