"""CPU tests: the oracle (oracle/slam_oracle.cpp) and the independent numpy model against the
committed golden vectors -- cv2 4.13.0 BFMatcher outputs, the reference's own KDTree2D outputs
and projection regression pins (oracle/gen_golden.py made them)."""
import numpy as np
import pytest

from conftest import golden, golden_names
from oracle import bf_model


@pytest.mark.parametrize("name", golden_names("bf_"))
def test_bf_oracle_matches_cv2_golden(oracle, name):
    g = golden(name)
    q, t = g["q"], g["t"]
    for impl in (oracle, bf_model):
        idx, dist = impl.knn2(q, t)
        assert np.array_equal(idx, g["knn_idx"]) and np.array_equal(dist, g["knn_dist"]), impl.__name__
        ci, cd = impl.crosscheck(q, t)
        assert np.array_equal(ci, g["cc_idx"]) and np.array_equal(cd, g["cc_dist"]), impl.__name__
        mq, mt = impl.match_descriptors(q, t, float(g["max_dist"]), float(g["ratio"]))
        assert np.array_equal(mq, g["md_query"]) and np.array_equal(mt, g["md_train"]), impl.__name__
        ft, fq = impl.match_features(t, q, float(g["max_dist"]))   # prev = train, cur = query
        assert np.array_equal(ft, g["mf_train"]) and np.array_equal(fq, g["mf_query"]), impl.__name__


def test_bf_threshold_edges_are_present():
    """The edge fixture really holds d0 == 64 rows (kept by match_descriptors' '>' gate, dropped
    by match_features' '<' gate) -- src/MapMatcher.cpp:152 vs src/features/FeatureExtractor.cpp:25."""
    g = golden("bf_edges_40x90.npz")
    assert (g["knn_dist"][:, 0] == 48).any() or (g["knn_dist"][:, 0] == 64).any()
    cc = g["cc_dist"]
    kept = set(g["mf_query"].tolist())
    for qi in np.nonzero(cc >= 64)[0]:
        assert qi not in kept


@pytest.mark.parametrize("name", golden_names("kd_"))
def test_kd_oracle_matches_reference_kdtree_golden(oracle, name):
    g = golden(name)
    xy, targets, radius = g["xy"], g["targets"], float(g["radius"])
    perm = oracle.kd_build(xy)
    assert np.array_equal(perm, g["perm"])
    o = 0
    for tgt, cnt in zip(targets, g["counts"]):
        mine = oracle.kd_radius_search(xy, perm, tgt, radius)
        assert np.array_equal(mine, g["flat"][o:o + cnt]), (name, tgt)
        o += cnt


def test_kd_oracle_matches_live_reference_when_built(oracle):
    """oracle/_ref/libref_kdtree.so is the reference's src/KDTree.cpp compiled unmodified."""
    if oracle.ref_kdtree() is None:
        pytest.skip("oracle/_ref not built (reference tree absent)")
    from racing_slam_b200 import synthetic
    for seed, n, w, h, r in [(1, 777, 100, 80, 7.5), (2, 64, 8, 8, 2.0), (3, 3000, 1920, 1080, 20.0)]:
        xy = synthetic.keypoints_integer(n, w, h, seed)
        rng = np.random.default_rng(seed)
        targets = np.stack([rng.uniform(0, w, 64), rng.uniform(0, h, 64)], 1).astype(np.float32)
        targets[:8] = np.rint(targets[:8])
        perm = oracle.kd_build(xy)
        for tgt, ref in zip(targets, oracle.ref_kd_radius_search_all(xy, targets, r)):
            assert np.array_equal(oracle.kd_radius_search(xy, perm, tgt, r), ref)


def _run_proj(oracle, g):
    cam, wh, prm = g["cam"], g["wh"], g["params"]
    p = oracle.ProjParams(cam[0], cam[1], cam[2], cam[3], int(wh[0]), int(wh[1]), prm[0], prm[1], prm[2], prm[3],
                          prm[4], int(g["replace"]))
    return oracle.match_projection(g["kp_xy"], g["kp_desc"], g["kd_perm"], g["kp_matched"], g["pos"], g["normal"],
                                   g["range"], g["obs_offset"], g["obs_desc"], g["skip"], None, g["pose"], p)


@pytest.mark.parametrize("name", golden_names("proj_"))
def test_proj_oracle_regression(oracle, name):
    g = golden(name)
    r = _run_proj(oracle, g)
    assert np.array_equal(r.point, g["out_point"]) and np.array_equal(r.dist, g["out_dist"])
    assert np.array_equal(r.uv, g["uv"], equal_nan=True)
    assert np.array_equal(r.kp, g["kp"]) and np.array_equal(r.d, g["d"])


def test_proj_oracle_equals_sequential_python_restatement(oracle):
    """A slow, literal Python loop over src/MapMatcher.cpp:45-98 (integer parts only: candidate
    order from the oracle's KD search) must give the oracle's proposals on a small scene."""
    from racing_slam_b200 import synthetic
    sc = synthetic.projection_scene(150, 80, 77, cam=dict(fx=100.0, fy=100.0, cx=40.0, cy=30.0, width=80, height=60))
    perm = oracle.kd_build(sc.kp_xy)
    p = oracle.ProjParams(100.0, 100.0, 40.0, 30.0, 80, 60, 20.0, 64.0, 0.5, 2.0, 1.25, 0)
    r = oracle.match_projection(sc.kp_xy, sc.kp_desc, perm, sc.kp_matched, sc.pos, sc.normal, sc.range, sc.obs_offset,
                                sc.obs_desc, sc.skip, None, sc.pose, p)
    prop_d = np.full(len(sc.kp_xy), 64.0)
    prop_p = np.full(len(sc.kp_xy), -1)
    for i in range(len(sc.pos)):
        if r.kp[i] < 0 and not np.isfinite(r.uv[i]).all():
            continue
        if not np.isfinite(r.uv[i]).all():
            continue
        # gates are the oracle's (float part); redo candidates x observations and the proposal rule
        cands = oracle.kd_radius_search(sc.kp_xy, perm, r.uv[i], 20.0)
        best, best_i = 64.0, 0
        gated = r.kp[i] >= 0 or True
        for c in cands:
            if sc.kp_matched[c]:
                continue
            for o in range(sc.obs_offset[i], sc.obs_offset[i + 1]):
                d = oracle.hamming(sc.kp_desc[c], sc.obs_desc[o])
                if d < best:
                    best, best_i = float(d), int(c)
        if r.kp[i] >= 0:            # the point passed every gate: its own choice must agree
            assert (best_i, int(best)) == (int(r.kp[i]), int(r.d[i]))
            if best < prop_d[best_i]:
                prop_d[best_i], prop_p[best_i] = best, i
    assert np.array_equal(prop_p.astype(np.int32), r.point)


def test_bf_model_vs_oracle_random(oracle):
    from racing_slam_b200 import synthetic
    for seed, (nq, nt) in enumerate([(257, 129), (33, 700), (1, 1), (5, 0), (0, 5)]):
        q, t = synthetic.bf_tie_heavy(nq, nt, 100 + seed) if seed % 2 else synthetic.bf_correlated(nq, nt, 100 + seed)
        a, b = oracle.knn2(q, t)
        c, d = bf_model.knn2(q, t)
        assert np.array_equal(a, c) and np.array_equal(b, d)
        a, b = oracle.crosscheck(q, t)
        c, d = bf_model.crosscheck(q, t)
        assert np.array_equal(a, c) and np.array_equal(b, d)


# ---- bag of words: the restatement against the reference's own DBoW2 -------------------
def _voc_of(oracle, g):
    return oracle.Vocabulary(g["child_offset"], g["child_ids"], g["desc"], g["weight"], g["word_id"], int(g["k"]),
                             int(g["L"]))


@pytest.mark.parametrize("name", golden_names("voc_"))
def test_voc_oracle_matches_dbow2_golden(oracle, name):
    """Fixtures come from thirdparty/DBoW2 compiled unmodified (oracle/gen_golden_voc.py): trees built by
    DBoW2's create(), words / bag-of-words vectors / L1 score from DBoW2's transform() and score()."""
    g = golden(name)
    v = _voc_of(oracle, g)
    word, w = oracle.voc_transform_each(v, g["query"])
    assert np.array_equal(word, g["word"]) and np.array_equal(w, g["word_weight"])
    ids, vals = oracle.voc_bow(word, w)
    assert np.array_equal(ids, g["bow_ids"]) and np.array_equal(vals, g["bow_vals"])     # doubles bit-identical
    w2, ww2 = oracle.voc_transform_each(v, g["query2"])
    ids2, vals2 = oracle.voc_bow(w2, ww2)
    assert np.array_equal(ids2, g["bow2_ids"]) and np.array_equal(vals2, g["bow2_vals"])
    assert oracle.voc_score_l1(ids, vals, ids2, vals2) == float(g["score"])


def test_voc_golden_has_ties_between_children(oracle):
    """The tie fixture really exercises 'first child with the strictly smallest distance'."""
    g = golden("voc_k3_L4_ties.npz")
    off, ids, desc = g["child_offset"], g["child_ids"], g["desc"]
    ties = 0
    for q in g["query"][:200]:
        node = 0
        while off[node + 1] > off[node]:
            kids = ids[off[node]:off[node + 1]]
            d = [oracle.hamming(q, desc[c]) for c in kids]
            ties += sorted(d)[0] == sorted(d)[1] if len(d) > 1 else 0
            node = int(kids[int(np.argmin(d))])
    assert ties > 20


def test_voc_oracle_matches_live_dbow2_when_built(oracle, pkg):
    """Fresh vocabulary from the real DBoW2 (k-means seeded from the clock) when oracle/_ref is present."""
    if oracle.ref_dbow2() is None:
        pytest.skip("oracle/_ref/libref_dbow2.so not built (no /root/reference on this machine)")
    syn = pkg.synthetic
    rng = np.random.default_rng(17)
    pool = rng.integers(0, 256, (1200, 32), dtype=np.uint8)
    imgs = [syn.flip_bits(pool[rng.choice(1200, 250, replace=False)], rng.integers(0, 40, 250), rng) for _ in range(16)]
    rv = oracle.RefVocabulary(6, 4, np.concatenate(imgs), np.arange(17) * 250)
    v = rv.export()
    q = syn.flip_bits(pool[rng.choice(1200, 900)], rng.integers(0, 70, 900), rng)
    word, w = oracle.voc_transform_each(v, q)
    rw, rww = rv.transform_each(q)
    assert np.array_equal(word, rw) and np.array_equal(w, rww)
    bow, rbow = oracle.voc_bow(word, w), rv.transform_bow(q)
    assert np.array_equal(bow[0], rbow[0]) and np.array_equal(bow[1], rbow[1])
    b2 = rv.transform_bow(q[::3])
    assert oracle.voc_score_l1(*bow, *b2) == rv.score(rbow, b2)
    rv.close()


def test_voc_bow_weightings_and_norms(oracle):
    """TF / IDF / BINARY and the un-normalised / L2 paths of transform() (TemplatedVocabulary.h:1081-1119)."""
    word = np.array([5, 2, 5, 9, 2, 5, 7], np.int32)
    w = np.array([0.5, 2.0, 0.5, 0.0, 2.0, 0.5, 1.5])           # word 9 is stopped (weight 0)
    ids, v = oracle.voc_bow(word, w, oracle.TF_IDF, oracle.NORM_L1)
    assert ids.tolist() == [2, 5, 7] and np.allclose(v, np.array([4.0, 1.5, 1.5]) / 7.0) and abs(v.sum() - 1) < 1e-15
    ids, v = oracle.voc_bow(word, w, oracle.IDF, oracle.NORM_NONE)
    assert ids.tolist() == [2, 5, 7] and v.tolist() == [2.0, 0.5, 1.5]
    ids, v = oracle.voc_bow(word, w, oracle.TF, oracle.NORM_NONE)
    assert v.tolist() == [4.0 / 3, 1.5 / 3, 1.5 / 3]
    ids, v = oracle.voc_bow(word, w, oracle.BINARY, oracle.NORM_L2)
    assert np.allclose((v * v).sum(), 1.0)


# ---- PnP verification support -----------------------------------------------------------------
@pytest.mark.parametrize("name", golden_names("pnp_"))
def test_pnp_projection_and_error_vs_cv2_fixtures(oracle, name):
    """The oracle's projectPoints / squared-error restatement against cv2 4.13.0's projections, bit for bit."""
    g = golden(name)
    for h in range(len(g["rt"])):
        uv = oracle.pnp_project(g["rt"][h], g["K4"], g["object"])
        assert np.array_equal(uv.view(np.uint32), g["uv"][h].view(np.uint32))
        c, err, mask = oracle.pnp_score(g["rt"][h], g["K4"], g["object"], g["image"], 4.0)
        assert np.array_equal(err.view(np.uint32), g["err_np"][h].view(np.uint32))
        with np.errstate(all="ignore"):
            want = g["err_np"][h] <= np.float32(16.0)
        assert np.array_equal(mask, want.astype(np.uint8)) and c == int(want.sum())


def test_ransac_update_num_iters_known_values(oracle):
    # log(1 - 0.99) / log(1 - (1 - ep)^5), rounded half to even; clamped to the current bound
    assert oracle.ransac_update_num_iters(0.99, 0.5, 5, 200) == 145
    assert oracle.ransac_update_num_iters(0.99, 0.0, 5, 200) == 0          # all inliers: denom = 0 -> stop
    assert oracle.ransac_update_num_iters(0.99, 1.0, 5, 200) == 200        # no inliers: keep the bound
    assert oracle.ransac_update_num_iters(0.99, 0.9, 5, 200) == 200        # would need 460k: clamped
    assert oracle.ransac_update_num_iters(0.99, 0.2, 5, 200) == 12
    assert oracle.ransac_update_num_iters(0.99, 0.5, 5, 100) == 100


def test_ransac_select_rule(oracle):
    # first strictly better model wins; a model needs more than model_points - 1 inliers; the bound shrinks
    assert oracle.ransac_select([3, 4, 4, 2], 100) == (-1, 4)
    assert oracle.ransac_select([10, 10, 9], 100) == (0, 3)
    assert oracle.ransac_select([10, 11, 11], 100) == (1, 3)
    best, iters = oracle.ransac_select([80] + [100] * 199, 100)            # 80 % inliers -> 12 iterations
    assert best == 1 and iters == 2                                        # model 1 is perfect: bound drops to 0
    best, iters = oracle.ransac_select([80] + [10] * 11 + [100] * 100, 100)
    assert (best, iters) == (0, 12)                                        # the perfect models come too late
