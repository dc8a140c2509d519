"""slamgpu_knn2_dev_noalloc: the kNN-2 scan for callers that must not allocate (no counterpart in the reference).
Its split partials live in a fixed 8 MiB device array inside the
library, so it calls no cudaMalloc; nq and nt are limited to 8192.

Through the C ABI, against the existing entry point slamgpu_knn2_dev on the same device buffers: bit-exact (integer
indices and distances), two calls in a row on DIFFERENT inputs and sizes (the shared scratch keeps the previous
call's contents), then the first inputs again.  This file sorts last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _both(ctx, torch, q, t):
    nq, nt = len(q), len(t)
    dq, dt = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
    outs = []
    for call in (ctx.knn2_dev, ctx.knn2_dev_noalloc):
        idx = torch.full((max(nq, 1), 2), -7, dtype=torch.int32, device="cuda")
        dist = torch.full((max(nq, 1), 2), -7, dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        call(dq.data_ptr(), nq, dt.data_ptr(), nt, idx.data_ptr(), dist.data_ptr())
        ctx.sync()
        outs.append((idx.cpu().numpy()[:nq], dist.cpu().numpy()[:nq]))
    return outs


def test_no_alloc_knn2_equals_the_existing_entry_point_on_consecutive_calls(pkg):
    import torch
    syn = pkg.synthetic
    cases = [syn.bf_correlated(2000, 1500, 41),      # many splits of a short train set
             syn.bf_correlated(8192, 8000, 42),      # the size limit (slamgpu_knn2_dev runs this one on the tensor cores)
             syn.bf_correlated(300, 8192, 43),       # few query rows: the most splits per row
             syn.bf_correlated(2000, 1500, 41)]      # the first inputs again, over the later calls' leftovers
    cases.append((cases[0][0][:5], cases[0][1][:1]))  # one train row: every second neighbour missing
    with pkg.Context(0) as ctx:
        launches = ctx.launches
        for k, (q, t) in enumerate(cases):
            (ri, rd), (gi, gd) = _both(ctx, torch, q, t)
            assert np.array_equal(gi, ri) and np.array_equal(gd, rd), k
            assert (ri[:, 0] >= 0).all()
        assert ctx.launches > launches
        # no train rows: every neighbour missing, as the existing entry point reports it
        (ri, rd), (gi, gd) = _both(ctx, torch, cases[0][0][:64], cases[0][1][:0])
        assert np.array_equal(gi, ri) and np.array_equal(gd, rd) and (gi == -1).all()
        # beyond the fixed scratch: refused with a message, nothing launched, and the context stays usable
        big = np.zeros((8193, 32), np.uint8)
        dbig = torch.from_numpy(big).cuda()
        out = torch.empty((8193, 2), dtype=torch.int32, device="cuda")
        for nq, nt in ((8193, 100), (100, 8193)):
            with pytest.raises(pkg.SlamGpuError) as e:
                ctx.knn2_dev_noalloc(dbig.data_ptr(), nq, dbig.data_ptr(), nt, out.data_ptr(), out.data_ptr())
            assert e.value.code == pkg.binding.ERR_INVALID and "8192" in str(e.value)
        (ri, rd), (gi, gd) = _both(ctx, torch, *cases[0])
        assert np.array_equal(gi, ri) and np.array_equal(gd, rd)


def test_no_alloc_knn2_on_a_second_context_of_the_same_device(pkg):
    """The scratch belongs to the device, not to a context: a second context (its own stream) gets the same
    results once the first one is idle."""
    import torch
    q, t = pkg.synthetic.bf_correlated(1000, 3000, 44)
    with pkg.Context(0) as a, pkg.Context(0) as b:
        (ri, rd), (gi, gd) = _both(a, torch, q, t)
        assert np.array_equal(gi, ri) and np.array_equal(gd, rd)
        (ri2, rd2), (gi2, gd2) = _both(b, torch, q[::-1].copy(), t)
        assert np.array_equal(gi2, ri2) and np.array_equal(gd2, rd2)
