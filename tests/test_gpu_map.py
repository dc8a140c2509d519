"""GPU parity tests of the MUTABLE resident map (SURVEY 8f N1 / N2 "with incremental updates") through the C ABI.

A host model of the reference's Map (points in vector order, observations per point in an order the host is
free to change, key frames with camera centres) is mutated the way the reference mutates it -- associate /
disassociate (src/Map.cpp:97-124), fuse (:78-95), remove from the middle (:63-76, Mapper::cull_points
src/Mapper.cpp:396-430), bundle-adjusted positions and key-frame poses (src/Optimization.cpp:371,534,
src/Mapper.cpp:391), map growth (src/Map.cpp:20-61) -- and every mutation is sent to the device as a DELTA
(slamgpu_points_set_observations / _set_position / _remove / _append_observed / slamgpu_observer_centers_update /
_compact).  After every step the resident set must behave exactly like the oracle fed a freshly flattened map:
bit-identical normals and ranges, identical match_projection and track_match results."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _params(pkg, cam, radius=20.0, replace=0):
    return pkg.ProjParams(cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["width"], cam["height"], radius, 64.0,
                          0.5, 2.0, 1.25, replace)


def _oparams(oracle, p):
    return oracle.ProjParams(p.fx, p.fy, p.cx, p.cy, p.width, p.height, p.radius, p.max_dist, p.min_cos, p.nearer,
                             p.further, p.replace)


class HostMap:
    """What the adapter's mirror keeps: point -> slot, key frame -> observer id; and the reference-side truth."""

    def __init__(self, ctx, set_id, rng, reserve=(0, 0, 0)):
        self.ctx, self.set_id, self.rng = ctx, set_id, rng
        self.centers = {}            # observer id -> float32[3]
        self.points = []             # Map's vector: dicts {pos, obs: [(observer, desc)], slot}
        self.n_slots = 0
        ctx.points_create(set_id, *reserve)

    # ---- reference-side mutations, each mirrored by one delta call ----
    def set_centers(self, ids, centers):
        for k, c in zip(ids, centers):
            self.centers[int(k)] = np.asarray(c, np.float32)
        self.ctx.observer_centers_update(self.set_id, ids, centers)

    def _runs(self, pts):
        off = np.zeros(len(pts) + 1, np.int32)
        desc, kf = [], []
        for i, p in enumerate(pts):
            for k, d in p["obs"]:
                kf.append(k)
                desc.append(d)
            off[i + 1] = len(kf)
        desc = np.asarray(desc, np.uint8).reshape(-1, 32)
        return off, desc, np.asarray(kf, np.int32)

    def append(self, new_pts):
        for p in new_pts:
            p["slot"] = self.n_slots
            self.n_slots += 1
            self.points.append(p)
        off, desc, kf = self._runs(new_pts)
        self.ctx.points_append_observed(self.set_id, np.stack([p["pos"] for p in new_pts]), off, desc, kf)

    def restate_observations(self, pts):
        """after associate / disassociate / fuse: the whole run, in the host's (new) iteration order"""
        off, desc, kf = self._runs(pts)
        self.ctx.points_set_observations(self.set_id, [p["slot"] for p in pts], off, desc, kf)

    def set_positions(self, pts):
        self.ctx.points_set_position(self.set_id, [p["slot"] for p in pts], np.stack([p["pos"] for p in pts]))

    def remove(self, pts):
        ids = {id(p) for p in pts}
        self.points = [p for p in self.points if id(p) not in ids]        # vector::erase keeps the order
        self.ctx.points_remove(self.set_id, [p["slot"] for p in pts])

    def compact(self):
        remap, n = self.ctx.points_compact(self.set_id)
        assert n == len(self.points)
        for j, p in enumerate(self.points):
            assert remap[p["slot"]] == j                                   # order kept, tombstones gone
            p["slot"] = j
        assert (remap >= 0).sum() == n
        self.n_slots = n

    # ---- the freshly flattened map the oracle sees ----
    def flatten(self, oracle):
        pos = np.stack([p["pos"] for p in self.points]).astype(np.float32)
        off, desc, kf = self._runs(self.points)
        centers = np.stack([self.centers[int(k)] for k in kf]).astype(np.float32) if len(kf) else np.zeros((0, 3), np.float32)
        normal, rng_ = oracle.point_stats(pos, off, centers)
        slots = np.array([p["slot"] for p in self.points], np.int32)
        return pos, off, desc, normal, rng_, slots


def _check(hm, ctx, oracle, pkg, sc, p, tag):
    pos, off, desc, normal, rng_, slots = hm.flatten(oracle)
    n_slots, _ = ctx.points_count(hm.set_id)
    assert n_slots == hm.n_slots, tag
    live, used, garbage = ctx.points_live(hm.set_id)
    assert live == len(hm.points) and used - garbage == off[-1], tag
    # N2: normals and ranges re-derived on the device for the touched points are bit-identical to a full derivation
    g_normal, g_range = ctx.points_stats(hm.set_id, n_slots)
    assert np.array_equal(g_normal[slots].view(np.uint32), normal.view(np.uint32)), tag
    assert np.array_equal(g_range[slots].view(np.uint32), rng_.view(np.uint32)), tag
    # the matcher on the resident set == the oracle on the flattened map; slots keep map order, so ties resolve alike
    slot_to_live = np.full(n_slots + 1, -1, np.int32)
    slot_to_live[slots] = np.arange(len(slots), dtype=np.int32)
    perm = oracle.kd_build(sc.kp_xy)
    op = _oparams(oracle, p)
    rng = np.random.default_rng(len(tag) + len(slots))
    skip_live = (rng.random(len(slots)) < 0.05).astype(np.uint8)
    skip_slots = np.zeros(n_slots, np.uint8)
    skip_slots[slots] = skip_live
    ref = oracle.match_projection(sc.kp_xy, sc.kp_desc, perm, sc.kp_matched, pos, normal, rng_, off, desc, skip_live,
                                  None, sc.pose, op, want_stats=False)
    ctx.frame_upload(3, sc.kp_xy, sc.kp_desc, None, sc.kp_matched)
    gp, gd = ctx.match_projection(3, hm.set_id, len(sc.kp_xy), sc.pose, p, skip_slots, None)
    assert np.array_equal(slot_to_live[gp], ref.point) and np.array_equal(gd, ref.dist), tag
    assert (ref.point >= 0).sum() > 20, tag
    # the chained tracking pair with slot lists
    kf_live = np.sort(rng.choice(len(slots), len(slots) // 3, replace=False)).astype(np.int32)
    m_live = rng.choice(len(slots), len(slots) // 20, replace=False).astype(np.int32)
    pt = np.zeros(len(slots), np.uint8)
    pt[m_live] = 1
    in_kf = np.zeros(len(slots), np.uint8)
    in_kf[kf_live] = 1
    kpm = sc.kp_matched.copy()
    r1 = oracle.match_projection(sc.kp_xy, sc.kp_desc, perm, kpm, pos, normal, rng_, off, desc,
                                 (pt | (1 - in_kf)).astype(np.uint8), None, sc.pose, op, want_stats=False)
    kpm[r1.point >= 0] = 1
    pt[r1.point[r1.point >= 0]] = 1
    r2 = oracle.match_projection(sc.kp_xy, sc.kp_desc, perm, kpm, pos, normal, rng_, off, desc, pt, None, sc.pose, op,
                                 want_stats=False)
    kf_pt, _, map_pt, _ = ctx.track_match(3, hm.set_id, len(sc.kp_xy), sc.pose, p, slots[m_live], slots[kf_live])
    assert np.array_equal(slot_to_live[kf_pt], r1.point) and np.array_equal(slot_to_live[map_pt], r2.point), tag


def _make_points(sc, idx, rng, n_kf):
    pts = []
    for i in idx:
        n_obs = int(sc.obs_offset[i + 1] - sc.obs_offset[i])
        kfs = rng.choice(n_kf, min(n_obs, n_kf), replace=False)            # one observation per key frame
        pts.append(dict(pos=sc.pos[i].copy(), obs=[(int(k), sc.obs_desc[sc.obs_offset[i] + j].copy())
                                                  for j, k in enumerate(kfs)]))
    return pts


@pytest.mark.parametrize("n_points,nkp,seed,reserve", [
    (6000, 1500, 41, (0, 0, 0)),                 # minimum reserves: every kind of growth and an arena repack happen
    (20000, 2000, 42, (1 << 16, 1 << 18, 64)),   # a reserved map: no reallocation
])
def test_scripted_mutation_sequence_matches_a_fresh_flatten(oracle, pkg, n_points, nkp, seed, reserve):
    syn = pkg.synthetic
    sc = syn.projection_scene(n_points, nkp, seed)
    p = _params(pkg, syn.KITTI)
    rng = np.random.default_rng(seed)
    n_kf = 12
    with pkg.Context(0) as ctx:
        hm = HostMap(ctx, 5, rng, reserve)
        centers = rng.normal(0.0, 0.6, (n_kf, 3)).astype(np.float32)
        centers[-1] += 30.0                                                # a far observer: the gates reject a share
        hm.set_centers(np.arange(n_kf), centers)
        # the initial map arrives in three uneven blocks (map growth)
        first = n_points * 3 // 4
        for a, b in [(0, first // 2), (first // 2, first - 7), (first - 7, first)]:
            hm.append(_make_points(sc, range(a, b), rng, n_kf))
        _check(hm, ctx, oracle, pkg, sc, p, "initial")

        # Map::associate: points gain an observation (the host's unordered_map may reorder the run)
        pts = [hm.points[i] for i in rng.choice(len(hm.points), 300, replace=False)]
        for q in pts:
            free = [k for k in range(n_kf) if k not in {k2 for k2, _ in q["obs"]}]
            if free:
                q["obs"].append((int(rng.choice(free)), rng.integers(0, 256, 32, dtype=np.uint8)))
                rng.shuffle(q["obs"])
        hm.restate_observations(pts)
        _check(hm, ctx, oracle, pkg, sc, p, "associate")

        # Map::disassociate: points lose observations (some lose all of them: they can no longer match)
        pts = [hm.points[i] for i in rng.choice(len(hm.points), 300, replace=False)]
        for j, q in enumerate(pts):
            q["obs"] = [] if j % 50 == 0 else q["obs"][:max(1, len(q["obs"]) - 1)]
        hm.restate_observations(pts)
        _check(hm, ctx, oracle, pkg, sc, p, "disassociate")

        # Map::fuse: kept takes the observations of discarded it does not have; discarded is erased
        order = rng.choice(len(hm.points), 200, replace=False)
        kept, gone = [hm.points[i] for i in order[:100]], [hm.points[i] for i in order[100:]]
        for q, g in zip(kept, gone):
            have = {k for k, _ in q["obs"]}
            q["obs"] += [(k, d) for k, d in g["obs"] if k not in have]
        hm.restate_observations(kept)
        hm.remove(gone)
        _check(hm, ctx, oracle, pkg, sc, p, "fuse")

        # Mapper::cull_points: erase from the middle of the vector (later points move up: visit ranks change)
        hm.remove([hm.points[i] for i in rng.choice(len(hm.points) // 2, 400, replace=False)])
        _check(hm, ctx, oracle, pkg, sc, p, "cull")

        # bundle adjustment: positions rewritten, key-frame poses moved
        pts = [hm.points[i] for i in rng.choice(len(hm.points), 1000, replace=False)]
        for q in pts:
            q["pos"] = (q["pos"] + rng.normal(0, 0.02, 3)).astype(np.float32)
        hm.set_positions(pts)
        moved = np.array([1, 4, 7], np.int32)
        hm.set_centers(moved, np.stack([hm.centers[int(k)] for k in moved]) + rng.normal(0, 0.05, (3, 3)).astype(np.float32))
        _check(hm, ctx, oracle, pkg, sc, p, "bundle adjustment")

        # Mapper::insert: a new key frame observes new points and some old ones
        hm.set_centers([n_kf], rng.normal(0.0, 0.6, (1, 3)).astype(np.float32))
        new = _make_points(sc, range(first, n_points), rng, n_kf)
        for q in new[::2]:
            q["obs"].append((n_kf, rng.integers(0, 256, 32, dtype=np.uint8)))
        hm.append(new)
        pts = [hm.points[i] for i in rng.choice(len(hm.points) - len(new), 500, replace=False)]
        for q in pts:
            q["obs"].insert(0, (n_kf, rng.integers(0, 256, 32, dtype=np.uint8)))
        hm.restate_observations(pts)
        _check(hm, ctx, oracle, pkg, sc, p, "insert key frame")

        # tombstones squeezed out, order kept; then everything again on the renumbered map
        hm.compact()
        _check(hm, ctx, oracle, pkg, sc, p, "compact")
        hm.remove([hm.points[i] for i in rng.choice(len(hm.points), 50, replace=False)])
        pts = [hm.points[i] for i in rng.choice(len(hm.points), 2000, replace=False)]

        def add_obs(q):
            free = [k for k in range(n_kf + 1) if k not in {k2 for k2, _ in q["obs"]}]
            if free:
                q["obs"].append((int(rng.choice(free)), rng.integers(0, 256, 32, dtype=np.uint8)))

        for _ in range(4):                 # runs that grow are rewritten at the arena's top, over and over: the
            for q in pts:                  # garbage forces a repack (small reserve) or fits the reserve (large one)
                rng.shuffle(q["obs"])
                if len(q["obs"]) > 2:
                    q["obs"].pop()
                add_obs(q)
                add_obs(q)
            hm.restate_observations(pts)
        _check(hm, ctx, oracle, pkg, sc, p, "after compaction")
        live, used, garbage = ctx.points_live(5)
        if reserve == (0, 0, 0):
            assert garbage < used                  # the arena was squeezed / regrown along the way
        ctx.points_release(5)


def test_mutable_map_argument_errors(ctx, pkg):
    sc = pkg.synthetic.projection_scene(50, 30, 3)
    ctx.points_create(9)
    desc = sc.obs_desc[:2]
    with pytest.raises(pkg.SlamGpuError):          # observer without a centre
        ctx.points_append_observed(9, sc.pos[:1], [0, 2], desc, [0, 1])
    ctx.observer_centers_update(9, [0, 1], np.zeros((2, 3), np.float32))
    with pytest.raises(pkg.SlamGpuError):          # an observer twice in one call
        ctx.observer_centers_update(9, [1, 1], np.zeros((2, 3), np.float32))
    ctx.points_append_observed(9, sc.pos[:3], [0, 2, 2, 2], desc, [0, 1])
    with pytest.raises(pkg.SlamGpuError):          # the CSR entry points do not apply to a mutable set
        ctx.points_append(9, sc.pos[:1], sc.normal[:1], sc.range[:1], [0, 1], desc[:1])
    with pytest.raises(pkg.SlamGpuError):
        ctx.points_set_position(9, [3], sc.pos[:1])                       # outside
    with pytest.raises(pkg.SlamGpuError):
        ctx.points_set_position(9, [1, 1], sc.pos[:2])                    # twice
    ctx.points_remove(9, [1])
    with pytest.raises(pkg.SlamGpuError):
        ctx.points_set_position(9, [1], sc.pos[:1])                       # removed
    with pytest.raises(pkg.SlamGpuError):
        ctx.points_remove(9, [1])
    remap, n = ctx.points_compact(9)
    assert n == 2 and remap.tolist() == [0, -1, 1]
    ctx.points_upload(10, sc.pos, sc.normal, sc.range, sc.obs_offset, sc.obs_desc)
    with pytest.raises(pkg.SlamGpuError):          # and the delta calls do not apply to a CSR set
        ctx.points_remove(10, [0])
    ctx.points_release(9), ctx.points_release(10)
