"""GPU parity tests, brute-force family: libslamgpu through its C ABI against
the cv2 golden vectors and the CPU oracle.  Bit-exact: indices and integer distances."""
import numpy as np
import pytest

from conftest import golden, golden_names

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", params=["int", "tc", "auto"])
def ctx(request, pkg):
    """Every brute-force test runs on both engines and on the library's own choice: the integer XOR+POPC
    scan (bf_kernels.cu), the tensor-core scan (bf_tc_kernels.cu), and `auto` (by problem size).
    SLAMGPU_BF_ENGINE is read when the context is created."""
    import os
    old = os.environ.get("SLAMGPU_BF_ENGINE")
    if request.param == "auto":
        os.environ.pop("SLAMGPU_BF_ENGINE", None)
    else:
        os.environ["SLAMGPU_BF_ENGINE"] = request.param
    c = pkg.Context(0)
    if old is None:
        os.environ.pop("SLAMGPU_BF_ENGINE", None)
    else:
        os.environ["SLAMGPU_BF_ENGINE"] = old
    yield c
    c.close()


@pytest.mark.parametrize("name", golden_names("bf_"))
def test_golden_cv2(ctx, name):
    g = golden(name)
    q, t = g["q"], g["t"]
    idx, dist = ctx.knn2(q, t)
    assert np.array_equal(idx, g["knn_idx"]) and np.array_equal(dist, g["knn_dist"])
    ci, cd = ctx.crosscheck(q, t)
    assert np.array_equal(ci, g["cc_idx"]) and np.array_equal(cd, g["cc_dist"])
    mq, mt = ctx.match_descriptors(q, t, float(g["max_dist"]), float(g["ratio"]))
    assert np.array_equal(mq, g["md_query"]) and np.array_equal(mt, g["md_train"])
    ft, fq = ctx.match_features(t, q, float(g["max_dist"]))
    assert np.array_equal(ft, g["mf_train"]) and np.array_equal(fq, g["mf_query"])


SIZES = [(1, 1), (1, 2), (2, 1), (3, 5), (31, 33), (32, 32), (255, 257), (256, 128), (257, 129), (513, 1000),
         (1000, 513), (2000, 2000), (2049, 4097), (4000, 4000)]


@pytest.mark.parametrize("nq,nt", SIZES)
@pytest.mark.parametrize("kind", ["correlated", "ties"])
def test_knn2_and_crosscheck_vs_oracle(ctx, oracle, pkg, nq, nt, kind):
    gen = pkg.synthetic.bf_correlated if kind == "correlated" else pkg.synthetic.bf_tie_heavy
    q, t = gen(nq, nt, nq * 7 + nt)
    idx, dist = ctx.knn2(q, t)
    oi, od = oracle.knn2(q, t)
    assert np.array_equal(idx, oi) and np.array_equal(dist, od)
    ci, cd = ctx.crosscheck(q, t)
    oi, od = oracle.crosscheck(q, t)
    assert np.array_equal(ci, oi) and np.array_equal(cd, od)


@pytest.mark.parametrize("nq,nt", [(0, 0), (0, 9), (9, 0)])
def test_empty_sets(ctx, nq, nt):
    q = np.zeros((nq, 32), np.uint8)
    t = np.zeros((nt, 32), np.uint8)
    idx, dist = ctx.knn2(q, t)
    assert idx.shape == (nq, 2) and (idx == -1).all() and (dist == -1).all()
    ci, cd = ctx.crosscheck(q, t)
    assert (ci == -1).all() and (cd == -1).all()
    assert len(ctx.match_descriptors(q, t)[0]) == 0        # src/MapMatcher.cpp:139-141
    assert len(ctx.match_features(t, q)[0]) == 0


@pytest.mark.parametrize("max_dist,ratio", [(64.0, 0.75), (64.0, 0.8), (50.5, 0.6), (256.0, 1.0), (0.0, 0.75), (100.0, 0.5), (30.0, 0.9),
                                            (64.0, 1.5), (200.0, 0.8), (63.5, 0.7)])
def test_match_descriptors_filters_vs_oracle(ctx, oracle, pkg, max_dist, ratio):
    q, t = pkg.synthetic.bf_correlated(1500, 1100, 42, max_flip=90)
    mq, mt = ctx.match_descriptors(q, t, max_dist, ratio)
    oq, ot = oracle.match_descriptors(q, t, max_dist, ratio)
    assert np.array_equal(mq, oq) and np.array_equal(mt, ot)
    ft, fq = ctx.match_features(t, q, max_dist)
    o_t, o_q = oracle.match_features(t, q, max_dist)
    assert np.array_equal(ft, o_t) and np.array_equal(fq, o_q)


def test_match_descriptors_single_train_row_uses_k1(ctx, oracle, pkg):
    q, t = pkg.synthetic.bf_correlated(300, 1, 5, max_flip=20, frac_replaced=0.0)
    mq, mt = ctx.match_descriptors(q, t)
    oq, ot = oracle.match_descriptors(q, t)
    assert np.array_equal(mq, oq) and np.array_equal(mt, ot)


def test_full_size_8k_properties(ctx, oracle, pkg):
    """BASELINE C3 size (8000 x 8000): bit-exact against the oracle, plus size-independent
    properties -- match_features of a permuted copy recovers the permutation, and the result is
    invariant under swapping roles (mutuality)."""
    rng = np.random.default_rng(3)
    a = rng.integers(0, 256, (8000, 32), dtype=np.uint8)
    perm = rng.permutation(8000)
    b = pkg.synthetic.flip_bits(a[perm], rng.integers(0, 20, 8000), rng)
    tr, qu = ctx.match_features(a, b, 64.0)              # prev = a (train), cur = b (query)
    assert len(qu) == 8000 and np.array_equal(qu, np.arange(8000)) and np.array_equal(tr, perm)
    tr2, qu2 = ctx.match_features(b, a, 64.0)            # swapped roles: the same pairs
    assert {(int(x), int(y)) for x, y in zip(tr, qu)} == {(int(y), int(x)) for x, y in zip(tr2, qu2)}
    idx, dist = ctx.knn2(b, a)
    oi, od = oracle.knn2(b, a)
    assert np.array_equal(idx, oi) and np.array_equal(dist, od)
    assert (dist[:, 0] <= dist[:, 1]).all()


def test_c5_full_size_database_properties(ctx, oracle, pkg):
    """BASELINE C5 size (1000 key frames x 4000 descriptors, the 128 MB database bench.py uses): 2048 ordered
    pairs drawn over the whole triangle through the host-buffer batch call.  Size-independent properties: list
    structure, every reported match within the distance gate (popcount recomputed here), the two shards of a
    world of 2 reproduce the unsharded lists pair for pair, and a sample of pairs is bit-exact against the oracle."""
    db = pkg.synthetic.keyframe_db_fast(1000, 4000, 5)
    desc = db.reshape(-1, 32)
    off = (np.arange(1001, dtype=np.int64) * 4000).astype(np.int32)
    every = pkg.sharding.all_pairs(1000)
    assert len(every) == 499500
    rng = np.random.default_rng(55)
    pairs = every[np.sort(rng.choice(len(every), 2048, replace=False))]
    o, ql, tl = ctx.match_descriptors_batch(desc, off, pairs, 64.0, 0.75)
    cnt = np.diff(o)
    assert o[0] == 0 and o[-1] == len(ql) == len(tl) and (cnt >= 0).all() and (cnt <= 4000).all()
    assert cnt.sum() > 2048 * 100                       # the shared rows do match
    pair_of = np.repeat(np.arange(len(pairs)), cnt)
    inner = np.ones(len(ql), bool)
    inner[o[:-1][cnt > 0]] = False                      # first entry of every non-empty list
    assert (np.diff(ql)[inner[1:]] > 0).all()           # query rows strictly ascending within a pair
    assert ((ql >= 0) & (ql < 4000) & (tl >= 0) & (tl < 4000)).all()
    qa = db[pairs[pair_of, 0], ql].view(np.uint64)
    ta = db[pairs[pair_of, 1], tl].view(np.uint64)
    d = np.bitwise_count(qa ^ ta).sum(axis=1)
    assert d.max() <= 64
    # sharding: block-cyclic halves, concatenated back in pair order
    got = {}
    for r in range(2):
        mine = pkg.sharding.shard_pairs(pairs, r, 2, block=64)
        so, sq, st = ctx.match_descriptors_batch(desc, off, mine, 64.0, 0.75)
        for k, (fq, ft) in enumerate(mine):
            got[(int(fq), int(ft))] = (sq[so[k]:so[k + 1]], st[so[k]:so[k + 1]])
    assert len(got) == len(pairs)
    for k, (fq, ft) in enumerate(pairs):
        sq, st = got[(int(fq), int(ft))]
        assert np.array_equal(sq, ql[o[k]:o[k + 1]]) and np.array_equal(st, tl[o[k]:o[k + 1]]), k
    for k in rng.choice(len(pairs), 256, replace=False):     # the OpenMP oracle does a few hundred pairs a second
        oq, ot = oracle.match_descriptors(db[pairs[k, 0]], db[pairs[k, 1]], 64.0, 0.75)
        assert np.array_equal(ql[o[k]:o[k + 1]], oq) and np.array_equal(tl[o[k]:o[k + 1]], ot), k


def test_pinned_arrays_from_host_alloc_as_inputs_and_outputs(ctx, oracle, pkg):
    """slamgpu_host_alloc through the binding: a page-locked descriptor set and result arrays give the same lists."""
    db = pkg.synthetic.keyframe_db(4, 500, 21, shared_frac=0.4)
    desc = pkg.pinned_empty((4 * 500, 32), np.uint8)
    desc[...] = db.reshape(-1, 32)
    off = (np.arange(5) * 500).astype(np.int32)
    pairs = pkg.sharding.all_pairs(4)
    out = (pkg.pinned_empty(len(pairs) * 500, np.int32), pkg.pinned_empty(len(pairs) * 500, np.int32))
    o, ql, tl = ctx.match_descriptors_batch(desc, off, pairs, 64.0, 0.75, out=out)
    for k, (fq, ft) in enumerate(pairs):
        oq, ot = oracle.match_descriptors(db[fq], db[ft], 64.0, 0.75)
        assert np.array_equal(ql[o[k]:o[k + 1]], oq) and np.array_equal(tl[o[k]:o[k + 1]], ot), k
    del out, desc


def test_resident_keyframe_database_equals_the_batch_call(ctx, oracle, pkg):
    """slamgpu_kfdb_*: key frames appended one by one (ragged, one empty, growth across the first reservation of
    2^20 rows), then pair lists matched against the resident copy -- the same lists as the upload-per-call form and
    the oracle; new frames appended between calls are visible to the next call."""
    sizes = [300, 0, 257, 1, 640, 33, 512]
    db = pkg.synthetic.keyframe_db(len(sizes), 640, 9, shared_frac=0.4)
    frames = [db[i, :n] for i, n in enumerate(sizes)]
    kf = pkg.matcher.KeyFrameDatabase(ctx)
    for i, f in enumerate(frames[:5]):
        assert kf.add(f) == i
    assert ctx.kfdb_count(kf.db_id) == (5, sum(sizes[:5]))
    pairs = pkg.sharding.all_pairs(5)
    pairs = np.concatenate([pairs, pairs[:, ::-1]])
    o, ql, tl = kf.match_pairs(pairs)
    desc = np.concatenate(frames[:5])
    off = np.concatenate([[0], np.cumsum(sizes[:5])]).astype(np.int32)
    bo, bq, bt = ctx.match_descriptors_batch(desc, off, pairs, 64.0, 0.75)
    assert np.array_equal(o, bo) and np.array_equal(ql, bq) and np.array_equal(tl, bt)
    for f in frames[5:]:
        kf.add(f)
    big = pkg.synthetic.keyframe_db_fast(270, 4000, 31)          # 1.08 M rows: the row store grows and moves
    for f in big:
        kf.add(f)
    assert len(kf) == 7 + 270 and ctx.kfdb_count(kf.db_id)[1] == sum(sizes) + 270 * 4000
    pairs = np.array([[6, 4], [4, 6], [5, 0], [276, 7], [7, 276], [100, 99], [3, 6], [1, 2]], np.int32)
    o, ql, tl = kf.match_pairs(pairs, 64.0, 0.75)
    every = frames + list(big)
    for k, (fq, ft) in enumerate(pairs):
        oq, ot = oracle.match_descriptors(every[fq], every[ft], 64.0, 0.75)
        assert np.array_equal(ql[o[k]:o[k + 1]], oq) and np.array_equal(tl[o[k]:o[k + 1]], ot), (k, fq, ft)
    with pytest.raises(pkg.SlamGpuError):
        kf.match_pairs([[0, 277]])                                # unknown frame
    kf.close()
    with pytest.raises(pkg.SlamGpuError):
        ctx.kfdb_match_pairs(kf.db_id, [[0, 1]])                  # released


def test_batch_vs_single_and_oracle(ctx, oracle, pkg):
    """LoopDetector all-pairs form (BASELINE C5 shape, small): ragged frames, every ordered pair."""
    rng = np.random.default_rng(9)
    sizes = [300, 0, 257, 1, 640, 33]
    db = pkg.synthetic.keyframe_db(len(sizes), 640, 9, shared_frac=0.4)
    frames = [db[i, :n] for i, n in enumerate(sizes)]
    desc = np.concatenate(frames)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    pairs = pkg.sharding.all_pairs(len(sizes))
    pairs = np.concatenate([pairs, pairs[:, ::-1]])       # both directions
    o, ql, tl = ctx.match_descriptors_batch(desc, off, pairs, 64.0, 0.75)
    assert o[0] == 0 and o[-1] == len(ql) == len(tl)
    for k, (fq, ft) in enumerate(pairs):
        oq, ot = oracle.match_descriptors(frames[fq], frames[ft], 64.0, 0.75)
        assert np.array_equal(ql[o[k]:o[k + 1]], oq), (k, fq, ft)
        assert np.array_equal(tl[o[k]:o[k + 1]], ot), (k, fq, ft)
    assert rng is not None


@pytest.mark.parametrize("slice_pairs,pinned", [("7", False), ("1", False), ("4", True), ("0", True)])
def test_batch_in_slices_returns_the_same_lists(oracle, pkg, monkeypatch, slice_pairs, pinned):
    """The host-buffer batch call brings its lists back slice by slice while later slices are scanned
    (SLAMGPU_BATCH_SLICE forces small slices; 0 = the library's choice, one slice here); pinned output arrays
    receive the device copies directly.  Same lists either way, and an overflow in a middle slice is reported
    with the size needed."""
    import torch
    monkeypatch.setenv("SLAMGPU_BATCH_SLICE", slice_pairs)
    sizes = [300, 0, 257, 1, 640, 33]
    db = pkg.synthetic.keyframe_db(len(sizes), 640, 9, shared_frac=0.4)
    frames = [db[i, :n] for i, n in enumerate(sizes)]
    desc = np.concatenate(frames)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    pairs = pkg.sharding.all_pairs(len(sizes))
    pairs = np.concatenate([pairs, pairs[:, ::-1]])
    want_q, want_t, want_o = [], [], [0]
    for fq, ft in pairs:
        oq, ot = oracle.match_descriptors(frames[fq], frames[ft], 64.0, 0.75)
        want_q.append(oq), want_t.append(ot), want_o.append(want_o[-1] + len(oq))
    want_q, want_t = np.concatenate(want_q), np.concatenate(want_t)
    assert want_o[-1] > 200
    with pkg.Context(0) as c:
        out = None
        if pinned:
            out = (torch.zeros(4096, dtype=torch.int32).pin_memory().numpy(),
                   torch.zeros(4096, dtype=torch.int32).pin_memory().numpy())
        o, ql, tl = c.match_descriptors_batch(desc, off, pairs, 64.0, 0.75, out=out)
        assert np.array_equal(o, want_o) and np.array_equal(ql, want_q) and np.array_equal(tl, want_t)
        o2, ql2, tl2 = c.match_descriptors_batch(desc, off, pairs, 64.0, 0.75, capacity=want_o[-1])    # exact fit
        assert np.array_equal(o2, want_o) and np.array_equal(ql2, want_q) and np.array_equal(tl2, want_t)
        with pytest.raises(pkg.SlamGpuError) as e:
            c.match_descriptors_batch(desc, off, pairs, 64.0, 0.75, capacity=want_o[len(pairs) // 2])
        assert e.value.code == pkg.binding.ERR_NOMEM
        o3, ql3, _ = c.match_descriptors_batch(desc, off, pairs[:3], 64.0, 0.75)                        # still usable
        assert np.array_equal(o3, want_o[:4]) and np.array_equal(ql3, want_q[:want_o[3]])


@pytest.mark.parametrize("engine", ["int", "tc"])
def test_short_batch_into_pinned_lists(oracle, pkg, monkeypatch, engine):
    """The loop-query path (at most 256 pairs, pinned result arrays): one CTA per pair filters and compacts, a second
    launch places the blocks -- ragged and empty frames, frames longer than one 1024-row step, 1 / 5 / 256 pairs,
    repeated pairs, an undersized list reported with the size needed, both engines."""
    monkeypatch.setenv("SLAMGPU_BF_ENGINE", engine)
    sizes = [2500, 0, 1025, 1, 1024, 33, 3000, 700]
    db = pkg.synthetic.keyframe_db(len(sizes), 3000, 19, shared_frac=0.5)
    frames = [db[i, :n] for i, n in enumerate(sizes)]
    desc = np.concatenate(frames)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    rng = np.random.default_rng(4)
    with pkg.Context(0) as c:
        for n_pairs in (1, 5, 256):
            pairs = rng.integers(0, len(sizes), (n_pairs, 2)).astype(np.int32)
            if n_pairs == 5:
                pairs = np.array([[0, 6], [6, 0], [1, 0], [0, 1], [0, 6]], np.int32)
            want = [oracle.match_descriptors(frames[a], frames[b], 64.0, 0.75) for a, b in pairs]
            want_o = np.concatenate([[0], np.cumsum([len(w[0]) for w in want])])
            cap = int(want_o[-1]) + 7
            out = (pkg.pinned_empty(cap, np.int32), pkg.pinned_empty(cap, np.int32))
            for _ in range(2):                                    # twice: the scratch and the counters are reused
                o, ql, tl = c.match_descriptors_batch(desc, off, pairs, 64.0, 0.75, out=out)
                assert np.array_equal(o, want_o), n_pairs
                assert np.array_equal(ql, np.concatenate([w[0] for w in want])), n_pairs
                assert np.array_equal(tl, np.concatenate([w[1] for w in want])), n_pairs
            if want_o[-1] > 8:
                small = (pkg.pinned_empty(int(want_o[-1]) - 3, np.int32), pkg.pinned_empty(int(want_o[-1]) - 3, np.int32))
                with pytest.raises(pkg.SlamGpuError) as e:
                    c.match_descriptors_batch(desc, off, pairs, 64.0, 0.75, out=small)
                assert e.value.code == pkg.binding.ERR_NOMEM
                del small
            del out


@pytest.mark.parametrize("fuse,push", [("1", "1"), ("0", "1"), ("1", "0"), ("0", "0")])
def test_fp4_ragged_batches_fused_finish_and_pushed_lists(oracle, pkg, monkeypatch, fuse, push):
    """Batches on the FP4 engine with one train split per pair: the scan's epilogue merges, filters and writes the rows
    itself (SLAMGPU_TC4_FUSE_FINISH=0: separate finish kernel), and with pinned result lists the compaction and the
    lists' way home run on a second stream (SLAMGPU_BF_PUSH=0: host-queued copies).  Ragged frames -- empty, one
    row, one row over a 256-row tile, several tiles -- so that pairs have query tiles without rows (padding written as
    -1) and single-row train sets (k = 1: no ratio test); 300 pairs in slices of 64 (several slices in flight), the
    device-resident form (padded rows exposed to the caller), and an undersized pinned list."""
    import torch
    monkeypatch.setenv("SLAMGPU_BF_ENGINE", "tc")
    monkeypatch.setenv("SLAMGPU_TC4_FUSE_FINISH", fuse)
    monkeypatch.setenv("SLAMGPU_BF_PUSH", push)
    monkeypatch.setenv("SLAMGPU_BATCH_SLICE", "64")
    sizes = [900, 0, 257, 1, 256, 33, 1300, 700]
    db = pkg.synthetic.keyframe_db(len(sizes), 1300, 23, shared_frac=0.5)
    frames = [db[i, :n] for i, n in enumerate(sizes)]
    desc = np.concatenate(frames)
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    rng = np.random.default_rng(8)
    pairs = np.concatenate([pkg.sharding.all_pairs(len(sizes)), pkg.sharding.all_pairs(len(sizes))[:, ::-1],
                            rng.integers(0, len(sizes), (244, 2)).astype(np.int32)]).astype(np.int32)
    assert len(pairs) == 300
    want = {}
    for a, b in {(int(a), int(b)) for a, b in pairs}:
        want[(a, b)] = oracle.match_descriptors(frames[a], frames[b], 64.0, 0.75)
    want_q = np.concatenate([want[(int(a), int(b))][0] for a, b in pairs])
    want_t = np.concatenate([want[(int(a), int(b))][1] for a, b in pairs])
    want_o = np.concatenate([[0], np.cumsum([len(want[(int(a), int(b))][0]) for a, b in pairs])])
    assert want_o[-1] > 5000
    with pkg.Context(0) as c:
        for pinned in (True, False):
            out = None
            if pinned:
                out = (pkg.pinned_empty(int(want_o[-1]) + 5, np.int32), pkg.pinned_empty(int(want_o[-1]) + 5, np.int32))
            for _ in range(2):
                o, ql, tl = c.match_descriptors_batch(desc, off, pairs, 64.0, 0.75, out=out)
                assert np.array_equal(o, want_o) and np.array_equal(ql, want_q) and np.array_equal(tl, want_t), pinned
        small = (pkg.pinned_empty(int(want_o[150]), np.int32), pkg.pinned_empty(int(want_o[150]), np.int32))
        with pytest.raises(pkg.SlamGpuError) as e:
            c.match_descriptors_batch(desc, off, pairs, 64.0, 0.75, out=small)
        assert e.value.code == pkg.binding.ERR_NOMEM
        assert np.array_equal(small[0], want_q[:len(small[0])])          # what fits is in place
        # device-resident form: [pair][max_rows] rows, -1 where a pair has no such query row
        d_desc, d_off, d_pairs = torch.from_numpy(desc).cuda(), torch.from_numpy(off).cuda(), torch.from_numpy(pairs).cuda()
        max_rows = max(sizes)
        d_out = torch.full((len(pairs), max_rows), -9, dtype=torch.int32, device="cuda")
        d_cnt = torch.full((len(pairs),), -9, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        c.match_descriptors_batch_dev(d_desc.data_ptr(), d_off.data_ptr(), len(sizes), d_pairs.data_ptr(), len(pairs), max_rows,
                                      64.0, 0.75, d_out.data_ptr(), d_cnt.data_ptr())
        c.sync()
        rows, cnt = d_out.cpu().numpy(), d_cnt.cpu().numpy()
        for k, (a, b) in enumerate(pairs):
            wq, wt = want[(int(a), int(b))]
            expect = np.full(max_rows, -1, np.int32)
            expect[wq] = wt
            assert np.array_equal(rows[k], expect) and cnt[k] == len(wq), (k, a, b)


def test_batch_capacity_error(ctx, pkg):
    db = pkg.synthetic.keyframe_db(3, 128, 1, shared_frac=0.9, max_flip=5)
    desc = db.reshape(-1, 32)
    off = np.array([0, 128, 256, 384], np.int32)
    pairs = pkg.sharding.all_pairs(3)
    with pytest.raises(pkg.SlamGpuError) as e:
        ctx.match_descriptors_batch(desc, off, pairs, 64.0, 0.75, capacity=3)
    assert e.value.code == pkg.binding.ERR_NOMEM


def test_device_pointer_forms(ctx, oracle, pkg):
    import torch
    q, t = pkg.synthetic.bf_correlated(3000, 2500, 77)
    dq, dt = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
    idx = torch.empty((3000, 2), dtype=torch.int32, device="cuda")
    dist = torch.empty((3000, 2), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ctx.knn2_dev(dq.data_ptr(), 3000, dt.data_ptr(), 2500, idx.data_ptr(), dist.data_ptr())
    ctx.sync()
    oi, od = oracle.knn2(q, t)
    assert np.array_equal(idx.cpu().numpy(), oi) and np.array_equal(dist.cpu().numpy(), od)
    out_train = torch.empty(3000, dtype=torch.int32, device="cuda")
    count = torch.zeros(1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ctx.match_descriptors_dev(dq.data_ptr(), 3000, dt.data_ptr(), 2500, 64.0, 0.75, out_train.data_ptr(),
                              count.data_ptr())
    ctx.sync()
    oq, ot = oracle.match_descriptors(q, t)
    got = out_train.cpu().numpy()
    assert int(count.item()) == len(oq)
    assert np.array_equal(np.nonzero(got >= 0)[0], oq) and np.array_equal(got[got >= 0], ot)


def test_errors_are_loud(ctx, pkg):
    with pytest.raises(pkg.SlamGpuError):
        ctx.match_projection(12345, 0, 4, np.eye(4, dtype=np.float32).flatten(), pkg.ProjParams(), None, None)
    lib = pkg.binding.load()
    assert lib.slamgpu_knn2(ctx._h, None, 4, None, 4, None, None) == pkg.binding.ERR_INVALID


def test_popc_microbenchmark_runs(ctx):
    rate, n_sm, mhz = ctx.bench_popc()
    assert n_sm >= 100 and rate > 1e12 and 500 < mhz < 2500


@pytest.mark.parametrize("copies", [0.0, 0.001, 0.01, 0.08, 0.3, 1.0])
def test_fp4_engine_tie_heavy_4000x4000_at_every_hit_density(oracle, pkg, monkeypatch, copies):
    """The FP4 filtered scan decodes accumulators only where a threshold bit is set: a 4000 x 4000 pair forced onto
    the tensor-core engine with the share of query rows that have EXACT copies in the train set swept from none to
    all (every copy duplicated, so the kNN is full of ties at distance 0 and the ratio test sees d0 = d1 = 0), the
    rest drawn from a 24-row pool (ties at every distance).  Bit-exact against the oracle at every density."""
    monkeypatch.setenv("SLAMGPU_BF_ENGINE", "tc")
    rng = np.random.default_rng(int(copies * 1000) + 3)
    q, t = pkg.synthetic.bf_tie_heavy(4000, 4000, 19)
    q, t = q.copy(), t.copy()
    uni = rng.integers(0, 256, (4000, 32), dtype=np.uint8)         # uniform rows: no accidental matches
    half = rng.random(4000) < 0.5
    q[half] = uni[half]
    n = int(copies * 4000)
    if n:
        rows = rng.choice(4000, n, replace=False)
        dst = rng.choice(4000, n, replace=False)
        t[dst] = q[rows]                                           # exact copies ...
        t[rng.choice(4000, n // 2, replace=False)] = q[rows[:n // 2]]   # ... some of them twice: d0 = d1 = 0
    with pkg.Context(0) as c:
        mq, mt = c.match_descriptors(q, t, 64.0, 0.75)
        oq, ot = oracle.match_descriptors(q, t, 64.0, 0.75)
        assert np.array_equal(mq, oq) and np.array_equal(mt, ot)
        ft, fq = c.match_features(t, q, 64.0)
        o_t, o_q = oracle.match_features(t, q, 64.0)
        assert np.array_equal(ft, o_t) and np.array_equal(fq, o_q)
        for ratio in (0.5, 1.0):                                   # d0 > ratio * d1 with d0 = d1 = 0 keeps the row
            mq, mt = c.match_descriptors(q, t, 64.0, ratio)
            oq, ot = oracle.match_descriptors(q, t, 64.0, ratio)
            assert np.array_equal(mq, oq) and np.array_equal(mt, ot)
    if copies >= 0.08:
        assert len(oq) > 100


@pytest.mark.parametrize("one_scan", ["1", "0"])
def test_tensor_core_cross_check_in_one_scan_or_two(oracle, pkg, monkeypatch, one_scan):
    """match_features on the FP4 engine: by default ONE scan whose hits also compete for their train row's column
    minimum (atomicMin of distance << 22 | query row), SLAMGPU_TC4_CROSS_ONE_SCAN=0 = a row scan and a column scan.
    Rectangular, non-tile-sized sets full of ties (rows and columns with several partners at the same distance: the
    lowest index must win on both sides), every distance bound the engine takes (cut <= 250)."""
    monkeypatch.setenv("SLAMGPU_BF_ENGINE", "tc")
    monkeypatch.setenv("SLAMGPU_TC4_CROSS_ONE_SCAN", one_scan)
    rng = np.random.default_rng(77)
    matched = 0
    with pkg.Context(0) as c:
        for nprev, ncur in ((3001, 4999), (4999, 3001), (700, 4100)):
            prev, cur = pkg.synthetic.bf_tie_heavy(nprev, ncur, 5 + nprev)
            prev, cur = prev.copy(), cur.copy()
            k = min(nprev, ncur) // 3
            cur[rng.choice(ncur, k, replace=False)] = prev[rng.choice(nprev, k, replace=False)]     # exact copies
            cur[rng.choice(ncur, k, replace=False)] = prev[rng.choice(nprev, k, replace=True)]      # several copies of one row
            for max_dist in (1.0, 30.0, 64.0, 100.5, 200.0):
                ft, fq = c.match_features(prev, cur, max_dist)
                o_t, o_q = oracle.match_features(prev, cur, max_dist)
                assert np.array_equal(ft, o_t) and np.array_equal(fq, o_q), (nprev, ncur, max_dist)
                matched += len(o_q)
    assert matched > 300


def test_two_contexts_on_two_devices_in_one_process(oracle, pkg):
    """The opt-in to > 48 KB of dynamic shared memory is a per-device function attribute: a second context on
    another GPU must set it for itself (round 1 kept one process-wide flag).  Skipped on a one-GPU box."""
    if pkg.binding.device_count() < 2:
        pytest.skip("needs two GPUs in one process")
    q, t = pkg.synthetic.bf_correlated(4000, 4000, 5)
    oq, ot = oracle.match_descriptors(q, t, 64.0, 0.75)
    oi, od = oracle.knn2(q, t)
    with pkg.Context(0) as c0, pkg.Context(1) as c1:
        for c in (c0, c1, c0):
            mq, mt = c.match_descriptors(q, t, 64.0, 0.75)          # FP4 tensor-core scan (164 KB of shared memory)
            assert np.array_equal(mq, oq) and np.array_equal(mt, ot)
            idx, dist = c.knn2(q, t)                                # int8 tensor-core scan
            assert np.array_equal(idx, oi) and np.array_equal(dist, od)


def test_two_contexts_on_one_device(oracle, pkg):
    """Two contexts in one process (what the reference's three MapMatcher owners would hold), used alternately."""
    q, t = pkg.synthetic.bf_correlated(4000, 4000, 6)
    oq, ot = oracle.match_descriptors(q, t, 64.0, 0.75)
    with pkg.Context(0) as c0, pkg.Context(0) as c1:
        for c in (c0, c1, c0, c1):
            mq, mt = c.match_descriptors(q, t, 64.0, 0.75)
            assert np.array_equal(mq, oq) and np.array_equal(mt, ot)


@pytest.mark.parametrize("engine", ["int", "tc"])
def test_sizes_at_and_over_the_train_limit(pkg, monkeypatch, engine):
    """SLAMGPU_MAX_TRAIN rows (4 Mi: the 22 index bits of the packed key) is accepted -- the last train row wins
    when it is the only close one -- and one row more is refused."""
    monkeypatch.setenv("SLAMGPU_BF_ENGINE", engine)
    n = pkg.binding.MAX_TRAIN
    rng = np.random.default_rng(4)
    t = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    q = t[[n - 1, 0, n // 2, n - 1]].copy()
    q[3, 0] ^= 1                                                   # distance 1 from the last row
    with pkg.Context(0) as c:
        idx, dist = c.knn2(q, t)
        assert idx[:, 0].tolist() == [n - 1, 0, n // 2, n - 1] and dist[:, 0].tolist() == [0, 0, 0, 1]
        assert (dist[:, 1] > 60).all() and (idx[:, 1] >= 0).all() and (idx[:, 1] < n).all()
        with pytest.raises(pkg.SlamGpuError) as e:
            c.knn2(q, np.concatenate([t, t[:1]]))
        assert e.value.code == pkg.binding.ERR_INVALID


@pytest.mark.parametrize("nq,nt", [(4001, 3999), (2561, 4321), (4097, 513)])
def test_padding_rows_of_the_expanded_copy_are_never_matched(oracle, pkg, monkeypatch, nq, nt):
    """The tensor-core scans read whole tiles from the expanded copy: rows between a set's end and its pitch must
    behave as 'no row' whatever the scratch held before.  The scratch is poisoned with descriptors that would match
    (a larger call with exact copies) and then reused by ragged-size calls."""
    monkeypatch.setenv("SLAMGPU_BF_ENGINE", "tc")
    with pkg.Context(0) as c:
        big = np.random.default_rng(1).integers(0, 256, (9000, 32), dtype=np.uint8)
        c.match_descriptors(big, big, 64.0, 0.75)                  # leaves exact copies of `big` all over the scratch
        c.knn2(big, big)
        q = big[:nq].copy()
        t = np.random.default_rng(2).integers(0, 256, (nt, 32), dtype=np.uint8)   # unrelated to the poison
        t[::97] = q[: len(t[::97])]                                # a few true matches
        mq, mt = c.match_descriptors(q, t, 64.0, 0.75)
        oq, ot = oracle.match_descriptors(q, t, 64.0, 0.75)
        assert np.array_equal(mq, oq) and np.array_equal(mt, ot)
        idx, dist = c.knn2(q, t)
        oi, od = oracle.knn2(q, t)
        assert np.array_equal(idx, oi) and np.array_equal(dist, od)
        ft, fq = c.match_features(t, q, 64.0)
        o_t, o_q = oracle.match_features(t, q, 64.0)
        assert np.array_equal(ft, o_t) and np.array_equal(fq, o_q)


def test_debug_variant_variable_is_ignored_by_the_release_library(oracle, pkg, monkeypatch):
    """Round 1's library skipped the MMAs (bit 2), the TMEM reads (bit 3) or the rare path (bit 5) when
    SLAMGPU_TC_VARIANT said so; the release build must compute the same matches whatever the variable holds."""
    q, t = pkg.synthetic.bf_correlated(4000, 4000, 8)
    oq, ot = oracle.match_descriptors(q, t, 64.0, 0.75)
    oi, od = oracle.knn2(q, t)
    for v in ("4", "8", "32", "128", "1"):
        monkeypatch.setenv("SLAMGPU_TC_VARIANT", v)
        with pkg.Context(0) as c:
            mq, mt = c.match_descriptors(q, t, 64.0, 0.75)
            assert np.array_equal(mq, oq) and np.array_equal(mt, ot), v
            idx, dist = c.knn2(q, t)
            assert np.array_equal(idx, oi) and np.array_equal(dist, od), v
