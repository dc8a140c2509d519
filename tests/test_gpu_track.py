"""GPU parity tests of the per-frame tracking pair and of map growth through the C ABI.

slamgpu_track_match chains Tracker::match_with_last_key_frame and Tracker::match_with_map
(src/Tracker.cpp:232-248) on the device; the bar is the same as for the single calls: the results
must be bit-identical to the two reference calls made one after the other with Frame::add_map_match
(src/Frame.cpp:80-102) applied in between -- here the oracle called twice with the masks updated on
the host.  slamgpu_points_append must leave the resident set exactly as one full upload would."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _params(pkg, cam, radius=20.0, max_dist=64.0, replace=0):
    return pkg.ProjParams(cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["width"], cam["height"], radius, max_dist,
                          0.5, 2.0, 1.25, replace)


def _oparams(oracle, p):
    return oracle.ProjParams(p.fx, p.fy, p.cx, p.cy, p.width, p.height, p.radius, p.max_dist, p.min_cos, p.nearer,
                             p.further, p.replace)


def _oracle_track(oracle, sc, perm, kp_matched, matched_points, kf_points, op):
    """The two reference calls in sequence, with add_map_match between and after them."""
    n = len(sc.pos)
    pt_matched = np.zeros(n, np.uint8)
    pt_matched[matched_points] = 1
    kp_matched = kp_matched.copy()
    in_kf = np.zeros(n, np.uint8)
    in_kf[kf_points] = 1
    skip1 = (pt_matched | (1 - in_kf)).astype(np.uint8)          # :53 and :169
    r1 = oracle.match_projection(sc.kp_xy, sc.kp_desc, perm, kp_matched, sc.pos, sc.normal, sc.range, sc.obs_offset,
                                 sc.obs_desc, skip1, None, sc.pose, op, want_stats=False)
    for kp, pi in enumerate(r1.point):
        if pi >= 0:
            assert not kp_matched[kp] and not pt_matched[pi]
            kp_matched[kp] = 1
            pt_matched[pi] = 1
    r2 = oracle.match_projection(sc.kp_xy, sc.kp_desc, perm, kp_matched, sc.pos, sc.normal, sc.range, sc.obs_offset,
                                 sc.obs_desc, pt_matched, None, sc.pose, op, want_stats=False)
    for kp, pi in enumerate(r2.point):
        if pi >= 0:
            kp_matched[kp] = 1
            pt_matched[pi] = 1
    return r1, r2, kp_matched, pt_matched


TRACK_SCENES = [
    # n_points, nkp, seed, cam, n_kf, n_matched
    (10000, 2000, 2, "KITTI", 3000, 300),      # BASELINE C2 sizes
    (30000, 8000, 3, "HD1080", 3000, 1500),    # BASELINE C3 frame size, ~3k points seen by the key frame
    (2000, 500, 7, "KITTI", 0, 0),             # no key-frame points: stage 1 is empty
    (2000, 500, 8, "KITTI", 2000, 0),          # the key frame sees every point: stage 2 finds what is left
    (500, 300, 9, "KITTI", 100, 500),          # every point already matched in the frame: nothing at all
]


@pytest.mark.parametrize("n_points,nkp,seed,cam,n_kf,n_matched", TRACK_SCENES)
def test_track_match_equals_the_two_reference_calls(ctx, oracle, pkg, n_points, nkp, seed, cam, n_kf, n_matched):
    camd = getattr(pkg.synthetic, cam)
    sc = pkg.synthetic.projection_scene(n_points, nkp, seed, cam=camd)
    rng = np.random.default_rng(seed + 100)
    kf_points = np.sort(rng.choice(n_points, n_kf, replace=False)).astype(np.int32)
    matched_points = rng.choice(n_points, n_matched, replace=False).astype(np.int32)
    p = _params(pkg, camd)
    perm = oracle.kd_build(sc.kp_xy)
    r1, r2, kp_after, pt_after = _oracle_track(oracle, sc, perm, sc.kp_matched, matched_points, kf_points,
                                               _oparams(oracle, p))
    ctx.frame_upload(7, sc.kp_xy, sc.kp_desc, None, sc.kp_matched)
    ctx.points_upload(7, sc.pos, sc.normal, sc.range, sc.obs_offset, sc.obs_desc)
    kf_pt, kf_d, map_pt, map_d = ctx.track_match(7, 7, nkp, sc.pose, p, matched_points, kf_points)
    assert np.array_equal(kf_pt, r1.point) and np.array_equal(kf_d, r1.dist)
    assert np.array_equal(map_pt, r2.point) and np.array_equal(map_d, r2.dist)
    if n_matched < n_points and n_kf > 0:
        assert (kf_pt >= 0).sum() > 0
    # no keypoint and no point is handed out twice across the two stages
    both = np.concatenate([kf_pt[kf_pt >= 0], map_pt[map_pt >= 0]])
    assert len(np.unique(both)) == len(both) and not np.any((kf_pt >= 0) & (map_pt >= 0))
    # the resident frame now carries the host frame's flags after both add_map_match loops: a fuse-style
    # call that skips the matched points must agree with the oracle run on the updated masks
    ref3 = oracle.match_projection(sc.kp_xy, sc.kp_desc, perm, kp_after, sc.pos, sc.normal, sc.range, sc.obs_offset,
                                   sc.obs_desc, pt_after, None, sc.pose, _oparams(oracle, p), want_stats=False)
    op3, od3 = ctx.match_projection(7, 7, nkp, sc.pose, p, pt_after, None)
    assert np.array_equal(op3, ref3.point) and np.array_equal(od3, ref3.dist)
    # a second track call on the same frame starts from the flags the first one left behind
    mp2 = np.nonzero(pt_after)[0].astype(np.int32)
    q1, q2, _, _ = _oracle_track(oracle, sc, perm, kp_after, mp2, kf_points, _oparams(oracle, p))
    again = ctx.track_match(7, 7, nkp, sc.pose, p, mp2, kf_points)
    assert np.array_equal(again[0], q1.point) and np.array_equal(again[2], q2.point)
    ctx.frame_release(7), ctx.points_release(7)


def test_track_match_argument_errors(ctx, pkg):
    sc = pkg.synthetic.projection_scene(100, 50, 1)
    p = _params(pkg, pkg.synthetic.KITTI)
    ctx.frame_upload(8, sc.kp_xy, sc.kp_desc, None, None)
    ctx.points_upload(8, sc.pos, sc.normal, sc.range, sc.obs_offset, sc.obs_desc)
    with pytest.raises(pkg.SlamGpuError):
        ctx.track_match(8, 8, 50, sc.pose, p, [], [5, 3])            # not ascending
    with pytest.raises(pkg.SlamGpuError):
        ctx.track_match(8, 8, 50, sc.pose, p, [100], [])             # outside the set
    with pytest.raises(pkg.SlamGpuError):
        ctx.track_match(8, 99, 50, sc.pose, p, [], [])               # unknown set
    ctx.frame_release(8), ctx.points_release(8)


@pytest.mark.parametrize("derive,reserve", [(False, "0"), (True, "0"), (False, None)])
def test_points_append_equals_one_full_upload(oracle, pkg, monkeypatch, derive, reserve):
    """Map growth in blocks of uneven size (including growth past the doubled capacity and an empty
    initial set) leaves normals, ranges and match results bit-identical to one upload.  With
    SLAMGPU_MAP_RESERVE=0 every growth step is a reallocation with live data; by default the first growth
    reserves the whole run."""
    if reserve is not None:
        monkeypatch.setenv("SLAMGPU_MAP_RESERVE", reserve)
    with pkg.Context(0) as ctx:
        _append_equals_upload(ctx, oracle, pkg, derive)


def _append_equals_upload(ctx, oracle, pkg, derive):
    n, nkp = 9000, 1500
    sc = pkg.synthetic.projection_scene(n, nkp, 41)
    p = _params(pkg, pkg.synthetic.KITTI)
    nrm, rng_ = (None, None) if derive else (sc.normal, sc.range)
    cen = sc.obs_center if derive else None
    ctx.points_upload(20, sc.pos, nrm, rng_, sc.obs_offset, sc.obs_desc, cen)
    cuts = [0, 0, 1, 2, 700, 701, 5000, 5003, n]      # first block empty
    off = sc.obs_offset

    def block(a, b):
        o0, o1 = off[a], off[b]
        return (sc.pos[a:b], None if derive else sc.normal[a:b], None if derive else sc.range[a:b],
                (off[a:b + 1] - o0).astype(np.int32), sc.obs_desc[o0:o1], sc.obs_center[o0:o1] if derive else None)

    ctx.points_upload(21, *block(0, 0))
    for a, b in zip(cuts[:-1], cuts[1:]):
        ctx.points_append(21, *block(a, b))
    assert ctx.points_count(21) == (n, int(off[n])) == ctx.points_count(20)
    n_full, r_full = ctx.points_stats(20, n)
    n_app, r_app = ctx.points_stats(21, n)
    assert np.array_equal(n_full, n_app) and np.array_equal(r_full, r_app)
    if derive:
        on, orng = oracle.point_stats(sc.pos, sc.obs_offset, sc.obs_center)
        assert np.array_equal(n_app, on) and np.array_equal(r_app, orng)
    ctx.frame_upload(20, sc.kp_xy, sc.kp_desc, None, sc.kp_matched)
    a = ctx.match_projection(20, 20, nkp, sc.pose, p, sc.skip, None)
    b = ctx.match_projection(20, 21, nkp, sc.pose, p, sc.skip, None)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and (a[0] >= 0).sum() > 100
    ctx.frame_release(20), ctx.points_release(20), ctx.points_release(21)


def test_mirror_track_equals_the_reference_call_sequence(ctx, oracle, pkg):
    """MapMatcher.track() through the mirror classes == match_key_frame, add_map_match loop,
    match_map, add_map_match loop (src/Tracker.cpp:232-248) on an identical second world."""
    from test_gpu_proj import _build_world
    m = pkg.matcher
    cam, world_a, kfs_a, frame_a = _build_world(pkg, 77)
    _, world_b, kfs_b, frame_b = _build_world(pkg, 77)
    matcher = m.MapMatcher(cam, m.ORB_MAX_DISTANCE, m.NORM_HAMMING, ctx=ctx)
    idx_a = {id(p): i for i, p in enumerate(world_a)}
    idx_b = {id(p): i for i, p in enumerate(world_b)}
    kf_m, map_m = matcher.track(frame_a, world_a, kfs_a[2])
    want_kf = matcher.match_key_frame(frame_b, world_b, kfs_b[2])
    for x in want_kf:
        frame_b.add_map_match(x)
    want_map = matcher.match_map(frame_b, world_b)
    for x in want_map:
        frame_b.add_map_match(x)
    assert [(idx_a[id(x.point)], x.keypoint_index) for x in kf_m] == \
           [(idx_b[id(x.point)], x.keypoint_index) for x in want_kf]
    assert [(idx_a[id(x.point)], x.keypoint_index) for x in map_m] == \
           [(idx_b[id(x.point)], x.keypoint_index) for x in want_map]
    assert len(kf_m) > 5 and len(map_m) > 20
    assert frame_a.num_map_matches() == frame_b.num_map_matches()


def test_kd_build_on_the_library_thread_equals_the_inline_build(oracle, pkg, monkeypatch):
    """slamgpu_frame_upload hands the KD build to a host thread (default) or does it inline
    (SLAMGPU_KD_ASYNC=0): same permutation as the oracle, same matches, also when the frame is replaced
    or released while its build is still running."""
    sc = pkg.synthetic.projection_scene(8000, 8000, 61, cam=pkg.synthetic.HD1080)
    p = _params(pkg, pkg.synthetic.HD1080)
    perm = oracle.kd_build(sc.kp_xy)
    results = []
    for mode in ("1", "0"):
        monkeypatch.setenv("SLAMGPU_KD_ASYNC", mode)
        with pkg.Context(0) as c:
            c.points_upload(1, sc.pos, sc.normal, sc.range, sc.obs_offset, sc.obs_desc)
            for _ in range(3):                                    # replace the slot while a build may be in flight
                c.frame_upload(1, sc.kp_xy[::-1].copy(), sc.kp_desc, None, None)
            c.frame_upload(1, sc.kp_xy, sc.kp_desc, None, sc.kp_matched)
            c.frame_set_matched(1, sc.kp_matched)                  # does not need the tree
            results.append(c.match_projection(1, 1, 8000, sc.pose, p, sc.skip, None))
            assert np.array_equal(c.frame_kd_perm(1, 8000), perm)
            c.frame_upload(2, sc.kp_xy, sc.kp_desc, None, None)
            c.frame_release(2)                                     # released before anything joined the build
            c.frame_upload(3, sc.kp_xy, sc.kp_desc, None, None)    # left for the context's destructor to join
    ref = oracle.match_projection(sc.kp_xy, sc.kp_desc, perm, sc.kp_matched, sc.pos, sc.normal, sc.range,
                                  sc.obs_offset, sc.obs_desc, sc.skip, None, sc.pose, _oparams(oracle, p),
                                  want_stats=False)
    for op, od in results:
        assert np.array_equal(op, ref.point) and np.array_equal(od, ref.dist)


@pytest.mark.parametrize("two_pass", ["1", "2"])
def test_track_match_with_either_form_of_the_match_kernel(oracle, pkg, monkeypatch, two_pass):
    """The tracking pair with K4 forced to its gate + search form (1) and to the fused kernel (2): by default
    the key-frame subset takes the fused kernel and a large map the two passes."""
    monkeypatch.setenv("SLAMGPU_PROJ_TWOPASS", two_pass)
    with pkg.Context(0) as c:
        for scene in TRACK_SCENES[:2]:
            test_track_match_equals_the_two_reference_calls(c, oracle, pkg, *scene)


@pytest.mark.parametrize("grid_on_device", ["1", "0"])
@pytest.mark.parametrize("ties", [False, True])
def test_calls_that_arrive_before_the_kd_tree_speculate_and_repeat_on_ties(oracle, pkg, monkeypatch, ties, grid_on_device):
    """slamgpu_frame_upload builds the KD-tree on a host thread; a projection call that finds it missing runs on the
    tree-free grid and counts the ties only the KD visit order can settle (src/KDTree.cpp:52-82,
    src/MapMatcher.cpp:88).  SLAMGPU_KD_DELAY_MS makes the builder late on purpose.  Without ties the speculative
    results stand; with ties (low-entropy descriptors) the call waits for the tree, restores the frame's keypoint
    flags and repeats itself -- either way the results are the oracle's."""
    monkeypatch.setenv("SLAMGPU_KD_DELAY_MS", "40")
    monkeypatch.setenv("SLAMGPU_GRID_ON_DEVICE", grid_on_device)    # the tree-free grid from the device kernel / the host sort
    cam = dict(fx=300.0, fy=300.0, cx=160.0, cy=120.0, width=320, height=240)
    sc = pkg.synthetic.projection_scene(6000, 1500, 93, cam=cam)
    rng = np.random.default_rng(93)
    if ties:
        sc.kp_desc[:] = rng.integers(0, 2, sc.kp_desc.shape, dtype=np.uint8) * 255
        sc.obs_desc[:] = rng.integers(0, 2, sc.obs_desc.shape, dtype=np.uint8) * 255
    p = _params(pkg, cam, 20.0, 200.0 if ties else 64.0)
    op = _oparams(oracle, p)
    perm = oracle.kd_build(sc.kp_xy)
    kf_points = np.sort(rng.choice(6000, 2000, replace=False)).astype(np.int32)
    matched_points = rng.choice(6000, 300, replace=False).astype(np.int32)
    with pkg.Context(0) as c:
        c.points_upload(1, sc.pos, sc.normal, sc.range, sc.obs_offset, sc.obs_desc)
        # one projection call
        c.frame_upload(1, sc.kp_xy, sc.kp_desc, None, sc.kp_matched)
        gp, gd = c.match_projection(1, 1, 1500, sc.pose, p, sc.skip, None)
        ref = oracle.match_projection(sc.kp_xy, sc.kp_desc, perm, sc.kp_matched, sc.pos, sc.normal, sc.range, sc.obs_offset,
                                      sc.obs_desc, sc.skip, None, sc.pose, op, want_stats=False)
        assert np.array_equal(gp, ref.point) and np.array_equal(gd, ref.dist)
        calls, repeats = c.speculation_count()
        assert calls == 1 and repeats == (1 if ties else 0)
        # the chained tracking pair on a freshly uploaded frame (its tree is late again)
        c.frame_upload(1, sc.kp_xy, sc.kp_desc, None, sc.kp_matched)
        r1, r2, kp_after, pt_after = _oracle_track(oracle, sc, perm, sc.kp_matched, matched_points, kf_points, op)
        kf_pt, kf_d, map_pt, map_d = c.track_match(1, 1, 1500, sc.pose, p, matched_points, kf_points)
        assert np.array_equal(kf_pt, r1.point) and np.array_equal(kf_d, r1.dist)
        assert np.array_equal(map_pt, r2.point) and np.array_equal(map_d, r2.dist)
        calls, repeats = c.speculation_count()
        assert calls == 2 and repeats == (2 if ties else 0)
        # the frame's flags are those after both add_map_match loops (restored, then rewritten by the repeat)
        ref3 = oracle.match_projection(sc.kp_xy, sc.kp_desc, perm, kp_after, sc.pos, sc.normal, sc.range, sc.obs_offset,
                                       sc.obs_desc, pt_after, None, sc.pose, op, want_stats=False)
        op3, od3 = c.match_projection(1, 1, 1500, sc.pose, p, pt_after, None)
        assert np.array_equal(op3, ref3.point) and np.array_equal(od3, ref3.dist)
        # SLAMGPU_KD_SPECULATE=0 is the old behaviour: wait for the tree
    monkeypatch.setenv("SLAMGPU_KD_SPECULATE", "0")
    with pkg.Context(0) as c:
        c.points_upload(1, sc.pos, sc.normal, sc.range, sc.obs_offset, sc.obs_desc)
        c.frame_upload(1, sc.kp_xy, sc.kp_desc, None, sc.kp_matched)
        gp, gd = c.match_projection(1, 1, 1500, sc.pose, p, sc.skip, None)
        assert np.array_equal(gp, ref.point) and np.array_equal(gd, ref.dist) and c.speculation_count() == (0, 0)


@pytest.mark.parametrize("case", ["clustered", "one_cell", "wide", "huge_frame"])
def test_tree_free_grid_built_on_the_device_serves_speculative_calls(oracle, pkg, monkeypatch, case):
    """grid_build_kernel (one CTA, cells counted in shared memory) against the oracle through speculative calls:
    keypoints piled into a few cells, all in ONE cell, a sparse grid with more than 12k cells (dynamic shared memory
    above 48 KB), and a frame too large for the kernel (the host sort takes over).  The tree is held back so that the
    calls really run on this grid."""
    monkeypatch.setenv("SLAMGPU_KD_DELAY_MS", "60")
    rng = np.random.default_rng(7)
    cam = dict(pkg.synthetic.HD1080)
    n_points, nkp = 4000, 3000
    if case == "huge_frame":
        nkp = 70000
    sc = pkg.synthetic.projection_scene(n_points, nkp, 31, cam=cam)
    if case == "clustered":
        sc.kp_xy[:] = (sc.kp_xy // 160) * 160 + rng.uniform(0, 12, sc.kp_xy.shape).astype(np.float32)
    elif case == "one_cell":
        sc.kp_xy[:] = np.float32([900.0, 500.0]) + rng.uniform(0, 10, sc.kp_xy.shape).astype(np.float32)
    elif case == "wide":
        monkeypatch.setenv("SLAMGPU_GRID_CELL", "8")               # 240 x 135 = 32400 cells
    p = _params(pkg, cam, 20.0, 64.0)
    ref = oracle.match_projection(sc.kp_xy, sc.kp_desc, oracle.kd_build(sc.kp_xy), sc.kp_matched, sc.pos, sc.normal, sc.range,
                                  sc.obs_offset, sc.obs_desc, sc.skip, None, sc.pose, _oparams(oracle, p), want_stats=False)
    with pkg.Context(0) as c:
        c.points_upload(1, sc.pos, sc.normal, sc.range, sc.obs_offset, sc.obs_desc)
        for _ in range(2):
            c.frame_upload(1, sc.kp_xy, sc.kp_desc, None, sc.kp_matched)
            gp, gd = c.match_projection(1, 1, nkp, sc.pose, p, sc.skip, None)
            assert np.array_equal(gp, ref.point) and np.array_equal(gd, ref.dist)
        assert c.speculation_count()[0] == 2
