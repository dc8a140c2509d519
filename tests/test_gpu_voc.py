"""GPU parity tests of the bag-of-words path through the C ABI: DBoW2's
TemplatedVocabulary::transform (word lookup and bag-of-words vectors) and L1Scoring::score, as LoopDetector
uses them before the brute-force verification (src/LoopDetector.cpp:85-92,346-373).

Bar: word ids bit-exact; bag-of-words values and scores are doubles and must be BIT-IDENTICAL too (the sums
are done in DBoW2's order).  Checked against fixtures produced by the reference's own DBoW2
(tests/golden/voc_*.npz, oracle/gen_golden_voc.py) and against the oracle on larger synthetic trees."""
import numpy as np
import pytest

from conftest import golden, golden_names

pytestmark = pytest.mark.gpu


def _upload(ctx, vid, v):
    ctx.vocab_upload(vid, v["child_offset"], v["child_ids"], v["desc"], v["weight"], v["word_id"])


def _ovoc(oracle, v):
    return oracle.Vocabulary(np.ascontiguousarray(v["child_offset"], np.int32), np.ascontiguousarray(v["child_ids"], np.int32),
                             np.ascontiguousarray(v["desc"], np.uint8), np.ascontiguousarray(v["weight"], np.float64),
                             np.ascontiguousarray(v["word_id"], np.int32))


@pytest.mark.parametrize("name", golden_names("voc_"))
def test_golden_from_the_reference_dbow2(ctx, name):
    g = golden(name)
    _upload(ctx, 1, g)
    word, w = ctx.vocab_transform(1, g["query"])
    assert np.array_equal(word, g["word"]) and np.array_equal(w, g["word_weight"])
    (ids, vals), (ids2, vals2) = ctx.vocab_bow(1, np.concatenate([g["query"], g["query2"]]),
                                              [0, len(g["query"]), len(g["query"]) + len(g["query2"])])
    assert np.array_equal(ids, g["bow_ids"]) and np.array_equal(vals, g["bow_vals"])
    assert np.array_equal(ids2, g["bow2_ids"]) and np.array_equal(vals2, g["bow2_vals"])
    s = ctx.vocab_score_l1((ids, vals), [(ids2, vals2), (ids, vals), (ids2[:0], vals2[:0])])
    assert s[0] == float(g["score"])
    assert s[2] == 0.0 and (len(ids) == 0 or abs(s[1] - 1.0) < 1e-12)
    ctx.vocab_release(1)


TREES = [
    # k, L, seed, queries, low_entropy
    (10, 4, 1, 8000, False),      # 11k nodes, C3-sized image
    (10, 5, 2, 8000, False),      # 111k nodes
    (3, 7, 3, 3000, False),       # narrow tree: the 4-lane kernel
    (20, 2, 4, 2000, False),      # more children than lanes in a group
    (10, 3, 5, 4000, True),       # byte-valued descriptors: ties between children at every level
    (1, 5, 6, 100, False),        # a chain: one child per node
]


@pytest.mark.parametrize("k,L,seed,nq,low", TREES)
def test_transform_vs_oracle_on_synthetic_trees(ctx, oracle, pkg, k, L, seed, nq, low):
    syn = pkg.synthetic
    v = syn.random_vocabulary(k, L, seed)
    q = syn.vocabulary_queries(v, nq, seed + 50)
    if low:
        rng = np.random.default_rng(seed)
        v["desc"] = (rng.integers(0, 2, v["desc"].shape, dtype=np.uint8) * 255).astype(np.uint8)
        q = (rng.integers(0, 2, q.shape, dtype=np.uint8) * 255).astype(np.uint8)
    _upload(ctx, 2, v)
    word, w = ctx.vocab_transform(2, q)
    ow, oww = oracle.voc_transform_each(_ovoc(oracle, v), q)
    assert np.array_equal(word, ow) and np.array_equal(w, oww)
    assert len(np.unique(word)) > min(5, k ** L - 1) or k == 1
    ctx.vocab_release(2)


def test_transform_device_pointer_form(ctx, oracle, pkg):
    import torch
    syn = pkg.synthetic
    v = syn.random_vocabulary(10, 4, 9)
    q = syn.vocabulary_queries(v, 5000, 10)
    _upload(ctx, 3, v)
    dq = torch.from_numpy(q).cuda()
    dword = torch.full((5000,), -7, dtype=torch.int32, device="cuda")
    dw = torch.zeros(5000, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    ctx.vocab_transform_dev(3, dq.data_ptr(), 5000, dword.data_ptr(), dw.data_ptr())
    ctx.sync()
    ow, oww = oracle.voc_transform_each(_ovoc(oracle, v), q)
    assert np.array_equal(dword.cpu().numpy(), ow) and np.array_equal(dw.cpu().numpy(), oww)
    ctx.vocab_release(3)


@pytest.mark.parametrize("weighting,norm", [(0, 1), (0, 0), (1, 2), (2, 0), (3, 1)])
def test_bow_batch_vs_oracle(ctx, oracle, pkg, weighting, norm):
    """Images of very different sizes in one call: empty, one feature, the maximum, repeated descriptors."""
    syn = pkg.synthetic
    v = syn.random_vocabulary(10, 3, 21)
    v["weight"][np.nonzero(v["word_id"] >= 0)[0][::7]] = 0.0            # some stopped words
    sizes = [0, 1, 4000, 16384, 37, 0, 8000]
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    q = syn.vocabulary_queries(v, int(off[-1]), 22)
    q[off[4]:off[5]] = q[off[4]]                                          # one image of identical descriptors
    _upload(ctx, 4, v)
    got = ctx.vocab_bow(4, q, off, weighting, norm)
    ov = _ovoc(oracle, v)
    for i, (ids, vals) in enumerate(got):
        word, w = oracle.voc_transform_each(ov, q[off[i]:off[i + 1]])
        oi, ovl = oracle.voc_bow(word, w, weighting, norm)
        assert np.array_equal(ids, oi) and np.array_equal(vals, ovl), f"image {i}"
    assert len(got[3][0]) > 500 and len(got[0][0]) == 0
    ctx.vocab_release(4)


def test_score_l1_against_a_database_vs_oracle(ctx, oracle, pkg):
    """score_candidates (src/LoopDetector.cpp:346-373): one key frame's vector against every earlier one."""
    syn = pkg.synthetic
    v = syn.random_vocabulary(10, 4, 31)
    n_db, per = 300, 1500
    q_all = syn.vocabulary_queries(v, n_db * per, 32)
    _upload(ctx, 5, v)
    off = (np.arange(n_db + 1) * per).astype(np.int32)
    db = ctx.vocab_bow(5, q_all, off)
    db[17] = (db[17][0][:0], db[17][1][:0])                               # an empty vector in the middle
    query = ctx.vocab_bow(5, np.concatenate([q_all[:per // 2], q_all[5 * per:5 * per + per // 2]]))[0]
    scores = ctx.vocab_score_l1(query, db)
    want = np.array([oracle.voc_score_l1(*query, *d) for d in db])
    assert np.array_equal(scores, want)
    assert scores[17] == 0.0 and scores[0] > np.median(scores) and scores[5] > np.median(scores)
    assert np.array_equal(ctx.vocab_score_l1((query[0][:0], query[1][:0]), db), np.zeros(n_db))
    ctx.vocab_release(5)


def test_resident_database_equals_the_one_shot_call(ctx, oracle, pkg):
    """slamgpu_bowdb_*: key-frame vectors appended one by one (growth past the first capacity included),
    scored in place; same doubles as the host-buffer call and the oracle."""
    syn = pkg.synthetic
    v = syn.random_vocabulary(10, 3, 51)
    _upload(ctx, 8, v)
    n_db = 1300                                                           # > 1024: the offset table grows once
    sizes = np.random.default_rng(52).integers(0, 900, n_db)
    sizes[3] = 0
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    q_all = syn.vocabulary_queries(v, int(off[-1]), 53)
    db = ctx.vocab_bow(8, q_all, off)
    assert ctx.bowdb_count(30) == 0
    for vec in db:
        ctx.bowdb_append(30, vec)
    assert ctx.bowdb_count(30) == n_db
    query = ctx.vocab_bow(8, q_all[off[10]:off[12]])[0]
    got = ctx.bowdb_score_l1(30, query)
    assert np.array_equal(got, ctx.vocab_score_l1(query, db))
    assert np.array_equal(got, np.array([oracle.voc_score_l1(*query, *d) for d in db]))
    assert got[3] == 0.0 and got[10] > 0.3 and got[11] > 0.3
    with pytest.raises(pkg.SlamGpuError):
        ctx.bowdb_append(30, (np.array([5, 5], np.int32), np.array([0.5, 0.5])))
    ctx.bowdb_release(30), ctx.vocab_release(8)
    with pytest.raises(pkg.SlamGpuError):
        ctx.bowdb_score_l1(30, query)


def test_argument_errors(ctx, pkg):
    syn = pkg.synthetic
    v = syn.random_vocabulary(4, 2, 41)
    bad = dict(v)
    bad["child_ids"] = v["child_ids"].copy()
    bad["child_ids"][3] = bad["child_ids"][2]                             # a node listed as a child twice
    with pytest.raises(pkg.SlamGpuError):
        _upload(ctx, 6, bad)
    _upload(ctx, 6, v)
    with pytest.raises(pkg.SlamGpuError):
        ctx.vocab_bow(6, np.zeros((pkg.binding.BOW_MAX_FEATURES + 1, 32), np.uint8))
    with pytest.raises(pkg.SlamGpuError):
        ctx.vocab_transform(99, np.zeros((4, 32), np.uint8))
    with pytest.raises(pkg.SlamGpuError):
        ctx.vocab_score_l1((np.array([3, 2], np.int32), np.array([0.5, 0.5])),
                           [(np.array([2], np.int32), np.array([1.0]))])             # query ids not ascending
    ctx.vocab_release(6)
    # a root-only vocabulary is "empty": no words (TemplatedVocabulary.h:1069-1072)
    ctx.vocab_upload(7, [0, 0], np.zeros(0, np.int32), np.zeros((1, 32), np.uint8), [0.0], [-1])
    word, _ = ctx.vocab_transform(7, np.zeros((3, 32), np.uint8))
    assert word.tolist() == [-1, -1, -1]
    assert len(ctx.vocab_bow(7, np.zeros((3, 32), np.uint8))[0][0]) == 0
    ctx.vocab_release(7)


def test_mirror_vocabulary_and_database_read_like_the_reference(ctx, pkg):
    """bow_of + score_candidates through the mirror classes, on a DBoW2-built tree (golden fixture)."""
    m = pkg.matcher
    g = golden("voc_k10_L3.npz")
    voc = m.OrbVocabulary(ctx, g["child_offset"], g["child_ids"], g["desc"], g["weight"], g["word_id"])
    bow = voc.transform(g["query"])
    bow2 = voc.transform(g["query2"])
    assert np.array_equal(bow[0], g["bow_ids"]) and np.array_equal(bow[1], g["bow_vals"])
    assert voc.score(bow, bow2) == float(g["score"])
    assert len(voc.transform(np.zeros((0, 32), np.uint8))[0]) == 0
    db = m.BowDatabase(ctx)
    assert len(db.scores(bow)) == 0
    assert db.add(bow2) == 0 and db.add(bow) == 1
    s = db.scores(bow)
    assert s[0] == float(g["score"]) and abs(s[1] - 1.0) < 1e-12
    db.close(), voc.close()
