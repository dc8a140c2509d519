"""GPU parity tests of the PnP verification support through the C ABI: verify_pnp's input
assembly (src/LoopDetector.cpp:188-197) and the inlier scoring / model choice cv::solvePnPRansac runs per RANSAC
model (src/LoopDetector.cpp:203-214).

Bar: projections feed a float error and a threshold test, so the bar is BIT-EXACT -- inlier counts, masks and the
float squared errors (compared as bit patterns) must equal (i) the fixtures computed from cv2 4.13.0's own
projectPoints (tests/golden/pnp_*.npz, oracle/gen_golden_pnp.py) and (ii) the oracle on larger ragged batches."""
import numpy as np
import pytest

from conftest import golden, golden_names

pytestmark = pytest.mark.gpu

K4 = np.array([718.856, 718.856, 607.19, 185.22], np.float64)


def _problems(seed, sizes, n_models, noise=1.0, outliers=0.4):
    import racing_slam_b200 as pkg
    return pkg.synthetic.pnp_problems(seed, sizes, n_models, noise, outliers)


def _oracle_problem(oracle, obj, img, rt, thr=4.0, **kw):
    cnt = np.array([oracle.pnp_score(m, K4, obj, img, thr)[0] for m in rt], np.int32) if len(obj) else np.zeros(len(rt), np.int32)
    best = oracle.ransac_select(cnt, len(obj), **kw)[0] if len(obj) else -1
    if best >= 0:
        c, _, mask = oracle.pnp_score(rt[best], K4, obj, img, thr)
    else:
        c, mask = 0, np.zeros(len(obj), np.uint8)
    return cnt, best, c, mask


@pytest.mark.parametrize("name", golden_names("pnp_"))
def test_golden_from_cv2_project_points(ctx, oracle, name):
    import torch
    g = golden(name)
    obj, img, rt, k4 = g["object"], g["image"], g["rt"], g["K4"]
    with np.errstate(all="ignore"):
        want = (g["err_np"] <= np.float32(16.0))
    (cnt, best, inl, mask), = ctx.pnp_ransac_score([obj], [img], [rt], k4, 4.0)
    assert np.array_equal(cnt, want.sum(1))
    ob = oracle.ransac_select(cnt, len(obj))[0]
    assert best == ob
    if best >= 0:
        assert np.array_equal(mask, want[best].astype(np.uint8)) and inl == want[best].sum()
    # float squared errors of every model, bit for bit (NaN / inf included)
    d = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a, dt)).cuda()
    d_obj, d_img, d_rt = d(obj, np.float32), d(img, np.float32), d(rt, np.float64)
    d_begin, d_cnt = d([0], np.int32), d([len(obj)], np.int32)
    d_mask = torch.empty(len(obj), dtype=torch.uint8, device="cuda")
    d_err = torch.empty(len(obj), dtype=torch.float32, device="cuda")
    d_inl = torch.empty(1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    for h in range(len(rt)):
        d_sel = d([h], np.int32)
        torch.cuda.synchronize()
        ctx.pnp_mask_dev(d_obj.data_ptr(), d_img.data_ptr(), d_begin.data_ptr(), d_cnt.data_ptr(), 1, d_rt.data_ptr(),
                         d_sel.data_ptr(), k4, 4.0, d_mask.data_ptr(), d_err.data_ptr(), d_inl.data_ptr())
        ctx.sync()
        assert np.array_equal(d_err.cpu().numpy().view(np.uint32), g["err_np"][h].view(np.uint32)), (name, h)
        assert np.array_equal(d_mask.cpu().numpy(), want[h].astype(np.uint8))
        assert int(d_inl.cpu()[0]) == int(want[h].sum())


@pytest.mark.parametrize("seed,sizes,n_models", [
    (1, [300], [200]),                                     # one verify_pnp call: 200 iterations
    (2, [12, 0, 1, 33, 64, 65, 4000, 5, 31, 32], [7, 3, 4, 0, 200, 9, 12, 40, 1, 2]),   # ragged, empty, no models
    (3, [int(x) for x in np.random.default_rng(3).integers(12, 700, 150)], [200] * 150),
])
def test_batch_vs_oracle(ctx, oracle, seed, sizes, n_models):
    objs, imgs, models = _problems(seed, sizes, n_models)
    got = ctx.pnp_ransac_score(objs, imgs, models, K4, 4.0)
    for p, (cnt, best, inl, mask) in enumerate(got):
        ocnt, obest, oinl, omask = _oracle_problem(oracle, objs[p], imgs[p], models[p])
        assert np.array_equal(cnt, ocnt), p
        assert best == obest and inl == oinl, p
        assert np.array_equal(mask, omask), p


def test_other_thresholds_and_ransac_parameters(ctx, oracle):
    objs, imgs, models = _problems(11, [500, 40], [60, 60], noise=2.0)
    for thr, conf, mp, iters in [(1.0, 0.99, 5, 200), (8.0, 0.999, 4, 50), (0.0, 0.5, 5, 10), (3.5, 0.99, 6, 1)]:
        got = ctx.pnp_ransac_score(objs, imgs, models, K4, thr, conf, mp, iters)
        for p, (cnt, best, inl, mask) in enumerate(got):
            ocnt, obest, oinl, omask = _oracle_problem(oracle, objs[p], imgs[p], models[p], thr, model_points=mp,
                                                       confidence=conf, max_iters=iters)
            assert np.array_equal(cnt, ocnt) and best == obest and inl == oinl and np.array_equal(mask, omask)


def test_assemble_after_batched_match_descriptors(ctx, oracle, pkg):
    """match_descriptors over a pair list -> correspondences of every pair on the device, as
    src/LoopDetector.cpp:181-197 builds them on the host."""
    import torch
    syn = pkg.synthetic
    rng = np.random.default_rng(21)
    rows = [400, 0, 257, 512, 3, 300]
    off = np.concatenate([[0], np.cumsum(rows)]).astype(np.int32)
    base = rng.integers(0, 256, (600, 32), dtype=np.uint8)
    desc = np.concatenate([syn.flip_bits(base[rng.choice(600, r, replace=False)], rng.integers(0, 30, r), rng)
                           if r else np.zeros((0, 32), np.uint8) for r in rows])
    pos = rng.normal(0, 10, (off[-1], 3)).astype(np.float32)
    xy = rng.uniform(0, 1241, (off[-1], 2)).astype(np.float32)
    pairs = np.array([[0, 2], [2, 0], [3, 0], [1, 0], [0, 1], [4, 5], [5, 3], [3, 3]], np.int32)
    max_rows = 512
    d = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a, dt)).cuda()
    d_desc, d_off, d_pairs, d_pos, d_xy = d(desc, np.uint8), d(off, np.int32), d(pairs, np.int32), d(pos, np.float32), d(xy, np.float32)
    n_pairs = len(pairs)
    d_train = torch.full((n_pairs, max_rows), -7, dtype=torch.int32, device="cuda")
    d_mcount = torch.zeros(n_pairs, dtype=torch.int32, device="cuda")
    d_obj = torch.zeros((n_pairs, max_rows, 3), dtype=torch.float32, device="cuda")
    d_img = torch.zeros((n_pairs, max_rows, 2), dtype=torch.float32, device="cuda")
    d_q = torch.zeros((n_pairs, max_rows), dtype=torch.int32, device="cuda")
    d_count = torch.zeros(n_pairs, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ctx.match_descriptors_batch_dev(d_desc.data_ptr(), d_off.data_ptr(), len(rows), d_pairs.data_ptr(), n_pairs, max_rows,
                                    64.0, 0.75, d_train.data_ptr(), d_mcount.data_ptr())
    ctx.pnp_assemble_batch_dev(d_train.data_ptr(), d_off.data_ptr(), len(rows), d_pairs.data_ptr(), n_pairs, max_rows,
                               d_pos.data_ptr(), d_xy.data_ptr(), d_obj.data_ptr(), d_img.data_ptr(), d_q.data_ptr(),
                               d_count.data_ptr())
    ctx.sync()
    count, mcount = d_count.cpu().numpy(), d_mcount.cpu().numpy()
    assert np.array_equal(count, mcount)
    assert count.sum() > 300
    obj, img, qq = d_obj.cpu().numpy(), d_img.cpu().numpy(), d_q.cpu().numpy()
    for p, (qf, tf) in enumerate(pairs):
        q, t = desc[off[qf]:off[qf + 1]], desc[off[tf]:off[tf + 1]]
        oq, ot = oracle.match_descriptors(q, t, 64.0, 0.75)
        mt = np.full(len(q), -1, np.int32)
        mt[oq] = ot
        oobj, oimg, oqi = oracle.pnp_assemble(mt, pos[off[tf]:off[tf + 1]], xy[off[qf]:off[qf + 1]])
        n = len(oqi)
        assert count[p] == n, p
        assert np.array_equal(qq[p, :n], oqi) and np.array_equal(obj[p, :n], oobj) and np.array_equal(img[p, :n], oimg), p


def test_full_size_properties(ctx):
    """LoopDetector-sized batch (2000 problems x 200 models): the two kernels agree (count of model h = inliers of
    its mask), counts are invariant under a permutation of the correspondences, and an exact model keeps every
    noise-free correspondence."""
    import torch
    rng = np.random.default_rng(5)
    sizes = [int(x) for x in rng.integers(20, 600, 2000)]
    objs, imgs, models = _problems(6, sizes, [200] * 2000, noise=0.0, outliers=0.3)
    got = ctx.pnp_ransac_score(objs, imgs, models, K4, 4.0)
    perm = [rng.permutation(n) for n in sizes]
    got2 = ctx.pnp_ransac_score([o[p] for o, p in zip(objs, perm)], [i[p] for i, p in zip(imgs, perm)], models, K4, 4.0)
    for p in range(0, 2000):
        cnt, best, inl, mask = got[p]
        cnt2, best2, inl2, mask2 = got2[p]
        assert np.array_equal(cnt, cnt2) and best == best2 and inl == inl2
        assert np.array_equal(mask[perm[p]], mask2)
        assert best >= 0 and inl == cnt[best] == int(mask.sum())
        n = sizes[p]                                   # model 3 is the exact pose: every non-outlier (70 %) is an inlier
        assert cnt[3] >= 0.7 * n - 4.5 * np.sqrt(0.21 * n) - 1
