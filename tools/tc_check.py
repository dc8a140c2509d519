#!/usr/bin/env python
"""GPU check of the tensor-core brute-force scan against the integer scan (both through the C ABI).

  python tools/tc_check.py                 # runs every variant in a subprocess with a timeout
  python tools/tc_check.py --variant 0     # one variant in this process
"""
import argparse
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def make_ctx(engine, variant=0):
    import racing_slam_b200 as pkg
    os.environ["SLAMGPU_BF_ENGINE"] = engine
    os.environ["SLAMGPU_TC_VARIANT"] = str(variant)
    return pkg.Context(0)


def compare(name, a, b):
    ia, da = a
    ib, db = b
    bad = np.nonzero((ia != ib).any(axis=1) | (da != db).any(axis=1))[0]
    print(f"  {name}: rows {len(ia)}  mismatching rows {len(bad)}", flush=True)
    for r in bad[:4]:
        print(f"    row {r}: int idx {ia[r]} dist {da[r]}   tc idx {ib[r]} dist {db[r]}")
    return len(bad)


def run_variant(variant, quick):
    import torch

    import racing_slam_b200 as pkg
    syn = pkg.synthetic
    ci = make_ctx("int")
    ct = make_ctx("tc", variant)
    total_bad = 0
    sizes = [(128, 256), (128, 512), (100, 300), (300, 257), (1, 33), (4000, 4000), (1000, 5000), (8000, 8000)]
    if quick:
        sizes = sizes[:4]
    for nq, nt in sizes:
        for kind, gen in (("correlated", syn.bf_correlated), ("ties", syn.bf_tie_heavy)):
            q, t = gen(nq, nt, 7 + nq + nt)
            total_bad += compare(f"knn2 {kind} {nq}x{nt}", ci.knn2(q, t), ct.knn2(q, t))
    if total_bad == 0 and not quick:
        q, t = syn.bf_correlated(3000, 2500, 11)
        a, b = ci.match_descriptors(q, t), ct.match_descriptors(q, t)
        ok = all(np.array_equal(x, y) for x, y in zip(a, b))
        print("  match_descriptors 3000x2500 equal:", ok, flush=True)
        total_bad += 0 if ok else 1
        for gen, gname in ((syn.bf_correlated, "correlated"), (syn.bf_tie_heavy, "ties")):
            q2, t2 = gen(2100, 1900, 21)
            for md, ra in ((64.0, 0.75), (30.0, 0.9), (100.0, 0.5), (200.0, 0.8), (64.0, 1.5), (0.0, 0.75), (63.5, 0.7)):
                a, b = ci.match_descriptors(q2, t2, md, ra), ct.match_descriptors(q2, t2, md, ra)
                ok = all(np.array_equal(x, y) for x, y in zip(a, b))
                if not ok:
                    print(f"  match_descriptors {gname} max_dist {md} ratio {ra}: MISMATCH ({len(a[0])} vs {len(b[0])})", flush=True)
                    total_bad += 1
            for md in (64.0, 20.0, 63.5, 1.0, 300.0):
                a, b = ci.match_features(t2, q2, md), ct.match_features(t2, q2, md)
                ok = all(np.array_equal(x, y) for x, y in zip(a, b))
                if not ok:
                    print(f"  match_features {gname} max_dist {md}: MISMATCH ({len(a[0])} vs {len(b[0])})", flush=True)
                    total_bad += 1
        print("  filter sweeps done, bad so far:", total_bad, flush=True)
        a, b = ci.crosscheck(q, t), ct.crosscheck(q, t)
        ok = all(np.array_equal(x, y) for x, y in zip(a, b))
        print("  crosscheck 3000x2500 equal:", ok, flush=True)
        total_bad += 0 if ok else 1
        # batched all-pairs on a small ragged database
        rng = np.random.default_rng(3)
        rows = [700, 1000, 333, 1024, 129]
        db = np.concatenate([syn.bf_correlated(r, r, 40 + i)[0] for i, r in enumerate(rows)])
        off = np.concatenate([[0], np.cumsum(rows)]).astype(np.int32)
        pairs = np.array([(i, j) for i in range(len(rows)) for j in range(len(rows)) if i != j], np.int32)
        a = ci.match_descriptors_batch(db, off, pairs)
        b = ct.match_descriptors_batch(db, off, pairs)
        ok = all(np.array_equal(x, y) for x, y in zip(a, b))
        print("  match_descriptors_batch ragged 5 frames equal:", ok, flush=True)
        total_bad += 0 if ok else 1
    if total_bad == 0 and not quick:
        # timing: 8000 x 8000 device-resident kNN-2 and the C5-shaped batch
        q, t = syn.bf_correlated(8000, 8000, 5)
        dq, dt = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
        di = torch.empty((8000, 2), dtype=torch.int32, device="cuda")
        dd = torch.empty((8000, 2), dtype=torch.int32, device="cuda")
        for name, ctx in (("int", ci), ("tc", ct)):
            st = torch.cuda.ExternalStream(ctx.stream)
            for _ in range(3):
                ctx.knn2_dev(dq.data_ptr(), 8000, dt.data_ptr(), 8000, di.data_ptr(), dd.data_ptr())
            ctx.sync()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(10):
                ctx.knn2_dev(dq.data_ptr(), 8000, dt.data_ptr(), 8000, di.data_ptr(), dd.data_ptr())
            e1.record(st)
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 100.0
            print(f"  knn2_dev 8000x8000 [{name}]: {us:.1f} us/call  {64e6 / us / 1e3:.1f} Gcmp/s", flush=True)
        nf, nd, npairs = 1000, 4000, 2048
        db = syn.keyframe_db_fast(nf, nd, 5)
        off = (np.arange(nf + 1, dtype=np.int64) * nd).astype(np.int32)
        pairs = pkg.sharding.all_pairs(nf)[:npairs].astype(np.int32)
        d_db = torch.from_numpy(db.reshape(-1, 32)).cuda()
        d_off = torch.from_numpy(off).cuda()
        d_pairs = torch.from_numpy(np.ascontiguousarray(pairs)).cuda()
        outs = {}
        for name, ctx in (("int", ci), ("tc", ct)):
            d_out = torch.empty((npairs, nd), dtype=torch.int32, device="cuda")
            d_cnt = torch.empty(npairs, dtype=torch.int32, device="cuda")
            st = torch.cuda.ExternalStream(ctx.stream)
            call = lambda: ctx.match_descriptors_batch_dev(d_db.data_ptr(), d_off.data_ptr(), nf, d_pairs.data_ptr(),  # noqa: E731
                                                           npairs, nd, 64.0, 0.75, d_out.data_ptr(), d_cnt.data_ptr())
            call()
            ctx.sync()
            ctx.profile_enable(True)
            ctx.profile_read(0)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            for _ in range(3):
                call()
            e1.record(st)
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 3
            scan_ms, scan_n = ctx.profile_read(0)
            ctx.profile_enable(False)
            print(f"  batch 2048 pairs of 4000x4000 [{name}]: {ms:.2f} ms/call  {npairs * 16e6 / ms / 1e6:.1f} Gcmp/s   "
                  f"scan kernel {scan_ms / max(scan_n, 1):.2f} ms x {scan_n}", flush=True)
            outs[name] = (d_out.cpu().numpy(), d_cnt.cpu().numpy())
        ok = np.array_equal(outs["int"][0], outs["tc"][0]) and np.array_equal(outs["int"][1], outs["tc"][1])
        print("  batch outputs equal:", ok, " matches:", int(outs["int"][1].sum()), flush=True)
        total_bad += 0 if ok else 1
    print(f"variant {variant}: total mismatches {total_bad}", flush=True)
    return total_bad


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--variant", type=int, default=None)
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--variants", default="0,2")
    args = ap.parse_args()
    if args.variant is not None:
        sys.exit(1 if run_variant(args.variant, args.quick) else 0)
    for v in [int(x) for x in args.variants.split(",")]:
        print(f"=== variant {v} (quick)", flush=True)
        t0 = time.time()
        try:
            r = subprocess.run([sys.executable, __file__, "--variant", str(v), "--quick"], timeout=150)
            print(f"=== variant {v}: exit {r.returncode} in {time.time() - t0:.0f} s", flush=True)
            if r.returncode != 0:
                continue
        except subprocess.TimeoutExpired:
            print(f"=== variant {v}: TIMEOUT", flush=True)
            continue
        print(f"=== variant {v} (full)", flush=True)
        try:
            r = subprocess.run([sys.executable, __file__, "--variant", str(v)], timeout=400)
            print(f"=== variant {v} full: exit {r.returncode}", flush=True)
        except subprocess.TimeoutExpired:
            print(f"=== variant {v} full: TIMEOUT", flush=True)


if __name__ == "__main__":
    main()
