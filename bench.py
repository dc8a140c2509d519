#!/usr/bin/env python
"""bench.py -- Hamming comparisons/s of the brute-force descriptor matcher on the LoopDetector
all-pairs workload (BASELINE.json configs[4] shape: 1000 key frames x 4000 ORB descriptors,
ordered pairs query > candidate, kNN-2 + max-distance 64 + Lowe ratio 0.75 per pair), sharded
by frame pair over N GPUs with no data-path collective.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

A step = one batch of `--pairs-per-step` frame pairs per rank (each 4000 x 4000 = 16 M
comparisons).  `value` times the device-resident C-ABI call (inputs already in HBM) with CUDA
events on the library's stream; `e2e` times the host-buffer C-ABI call
(slamgpu_match_descriptors_batch: pinned host descriptors + pair list in, compacted match lists
out, copies inside the timed region).  `--impl reference` times the reference's own CPU path for
the same pairs: cv2.BFMatcher(NORM_HAMMING).knnMatch(k=2) + the filters of
src/MapMatcher.cpp:151-161 (the very library call the reference makes), else the oracle port.
Prints ONE JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_FRAMES, N_DESC = 1000, 4000          # BASELINE.json configs[4]
MAX_DIST, RATIO = 64.0, 0.75           # src/features/OrbFeatureExtractor.h:14-17, src/MapMatcher.cpp:18
METRIC, UNIT = "hamming_comparisons_per_s", "G/s"
POPC_PER_CLK_PER_SM = 16               # XU pipe; measured on the box by slamgpu_bench_popc (profiles/)
POPC_PER_COMPARISON = 8                # 256 bits = 8 POPC32: the plain XOR+POPC formulation
ALU_PER_CLK_PER_SM = 64                # LOP3 / VIMNMX lanes per clock per SM (ncu: 15 ALU ops/cmp = 45 % in v1)
# the v2 kernel as written: 8 LOP3 xor + 8 LOP3 carry-save (xor3 / majority) + 3 VIMNMX.U16x2 per
# two comparisons on the ALU pipe, and 4 POPC on the XU pipe -> the ALU pipe binds
ALU_PER_COMPARISON = 17.5
POPC_PER_COMPARISON_CSA = 4
NCU_SCAN_DRAM_BYTES_512_PAIRS = 7.2e6  # profiles/r01_ncu_full_prof_bf_scan_v2.txt
NCU_TC4_DRAM_BYTES_1024_PAIRS = 30.0e6  # profiles/r02_final_ncu_full_bf_tc4_scan.txt (27.5 MB read + 2.5 MB written: with the fused finish the scan
                                        # writes the filtered rows, no partial-key records; r02_ncu_full_bf_tc4_scan.txt before: 57.3 MB)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pairs-per-step", type=int, default=32768,
                    help="frame pairs per rank per step (32768 x 16 M comparisons = 60 ms on the tensor cores, so that "
                         "a 5-step timed region is long enough for nvidia-smi clock samples)")
    ap.add_argument("--ref-pairs-per-step", type=int, default=24, help="frame pairs per step of the CPU arm")
    ap.add_argument("--frames", type=int, default=N_FRAMES)
    ap.add_argument("--desc", type=int, default=N_DESC)
    ap.add_argument("--no-also", action="store_true", help="skip the secondary C1-C4 measurements")
    return ap.parse_args()


def dist_env():
    return (int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)),
            int(os.environ.get("WORLD_SIZE", 1)))


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(hbm_gbs=float(p["hbm_gbs"]), sm_max_mhz=float(p.get("sm_max_mhz", 1965.0)),
                    bf16_tflops=float(p.get("bf16_tflops", 1590.0)),
                    bf16_tflops_sustained=float(p.get("bf16_tflops_sustained", 1400.0)), source="measured")
    return dict(hbm_gbs=6650.0, sm_max_mhz=1965.0, bf16_tflops=1590.0, bf16_tflops_sustained=1400.0, source="fallback")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [x.strip() for x in line.split(",")]))

    def stop(self, t0: float, t1: float):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        rows = [r for t, r in self.rows if t0 - 0.05 <= t <= t1 + 0.05] or [r for _, r in self.rows]
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[0])), mx.append(float(r[1])), power.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def make_workload(args, rank: int, world: int):
    """The key-frame database (identical on every rank: same seed -- replication without a
    collective) and this rank's block-cyclic shard of the ordered pair list."""
    import racing_slam_b200 as pkg
    db = pkg.synthetic.keyframe_db_fast(args.frames, args.desc, 5)
    off = (np.arange(args.frames + 1, dtype=np.int64) * args.desc).astype(np.int32)
    pairs = pkg.sharding.shard_pairs(pkg.sharding.all_pairs(args.frames), rank, world, block=64)
    return db, off, pairs


def workload_name(args) -> str:
    return (f"loop-closure all-pairs verification (BASELINE configs[4] shape): {args.frames} key frames x "
            f"{args.desc} ORB descriptors, kNN-2 + max-distance {MAX_DIST:g} + ratio {RATIO:g} per ordered pair "
            f"(query > candidate)")


def config_of(args) -> dict:
    """The same object in both arms: what is measured.  How much of it one step of an arm takes is under `step`."""
    return {"workload": workload_name(args), "key_frames": args.frames, "descriptors_per_key_frame": args.desc,
            "max_distance": MAX_DIST, "ratio": RATIO, "pairs_in_the_whole_job": args.frames * (args.frames - 1) // 2}


def step_pairs(pairs: np.ndarray, step: int, n: int) -> np.ndarray:
    """The n pairs of step `step`: a sliding window over this rank's shard (wraps around)."""
    idx = (np.arange(n, dtype=np.int64) + step * n) % len(pairs)
    return np.ascontiguousarray(pairs[idx])


# ------------------------------------------------------------------------------------------------
# CPU arm: the reference's own path for one frame pair (src/MapMatcher.cpp:145-161)
# ------------------------------------------------------------------------------------------------
def cpu_matcher():
    """(kind, cores, fn, describe).  kind "reference": cv2.BFMatcher, the same OpenCV call the
    reference makes (its C++ cannot be built here: no Eigen/OpenCV headers -- DESIGN.md).  Falls back
    to the oracle port (OpenMP) when cv2 is missing -- the only place bench.py executes oracle/."""
    cores = os.cpu_count() or 1
    try:
        import cv2
        cv2.setNumThreads(cores)

        def run(q, t):
            knn = cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(q, t, k=2 if len(t) >= 2 else 1)
            n = 0
            for c in knn:                                    # the reference's filter loop, :151-161
                if not c or c[0].distance > MAX_DIST:
                    continue
                if len(c) > 1 and c[0].distance > RATIO * c[1].distance:
                    continue
                n += 1
            return n
        return "reference", cores, run, f"cv2 {cv2.__version__} BFMatcher.knnMatch(k=2) + ratio filter, {cores} threads"
    except ImportError:
        from oracle import cpu_oracle

        def run(q, t):
            return len(cpu_oracle.match_descriptors(q, t, MAX_DIST, RATIO)[0])
        return "port", cores, run, f"oracle/slam_oracle.cpp orc_match_descriptors, OpenMP {cores} threads"


def time_cpu_pairs(run, db, pairs, budget_s=None):
    """Wall time of `run` over the given pairs (stops early when budget_s is exceeded)."""
    t0 = time.perf_counter()
    done, matches = 0, 0
    for fq, ft in pairs:
        matches += run(db[fq], db[ft])
        done += 1
        if budget_s is not None and time.perf_counter() - t0 > budget_s:
            break
    return time.perf_counter() - t0, done, matches


def run_reference(args, rank, world):
    if rank != 0:
        return                                       # rank 0 alone runs the CPU arm
    import racing_slam_b200 as pkg
    db = pkg.synthetic.keyframe_db_fast(args.frames, args.desc, 5)
    pairs = pkg.sharding.all_pairs(args.frames)
    kind, cores, run, what = cpu_matcher()
    n = args.ref_pairs_per_step
    per_pair = float(args.desc) * args.desc
    for w in range(args.warmup):
        time_cpu_pairs(run, db, step_pairs(pairs, w, min(n, 4)))
    t = 0.0
    for s in range(args.steps):
        dt, done, _ = time_cpu_pairs(run, db, step_pairs(pairs, args.warmup + s, n))
        assert done == n
        t += dt
    value = n * args.steps * per_pair / t / 1e9
    sample = f"{n} frame pairs of {args.desc}x{args.desc} per step ({what})"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t / args.steps * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": config_of(args),
        "step": {"pairs_per_step": n, "comparisons_per_step": n * per_pair, "sharding": "cpu, rank 0 only"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def cpu_baseline(args, db, budget_s=12.0):
    """Bounded sample of the same workload on the host cores (rank 0, N=1 only)."""
    import racing_slam_b200 as pkg
    kind, cores, run, what = cpu_matcher()
    pairs = pkg.sharding.all_pairs(args.frames)
    time_cpu_pairs(run, db, pairs[:2])
    dt, done, _ = time_cpu_pairs(run, db, pairs[2:2 + 4096], budget_s=budget_s)
    per_pair = float(args.desc) * args.desc
    out = {"value": done * per_pair / dt / 1e9, "unit": UNIT, "cores": cores, "kind": kind,
           "sample": f"{done} frame pairs of {args.desc}x{args.desc} in {dt:.1f} s ({what})"}
    if kind == "reference":                     # the oracle port beside it, for transparency
        from oracle import cpu_oracle
        port = lambda q, t: len(cpu_oracle.match_descriptors(q, t, MAX_DIST, RATIO)[0])  # noqa: E731
        time_cpu_pairs(port, db, pairs[:1])
        dt2, done2, _ = time_cpu_pairs(port, db, pairs[2:2 + 4096], budget_s=budget_s / 2)
        out["port"] = {"value": done2 * per_pair / dt2 / 1e9, "unit": UNIT, "cores": cores,
                       "sample": f"{done2} pairs, oracle/slam_oracle.cpp with OpenMP"}
    return out


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def popc_peak_gcmp(n_sm: int, mhz: float) -> float:
    """SURVEY 8d roofline: XU pipe with 8 POPC32 per comparison."""
    return n_sm * POPC_PER_CLK_PER_SM * mhz * 1e6 / POPC_PER_COMPARISON / 1e9


def int_peak_gcmp(n_sm: int, mhz: float) -> float:
    """Integer-pipe roofline of the kernel as written: the slower of ALU (17.5 ops) and XU (4 POPC)."""
    alu = n_sm * ALU_PER_CLK_PER_SM * mhz * 1e6 / ALU_PER_COMPARISON / 1e9
    xu = n_sm * POPC_PER_CLK_PER_SM * mhz * 1e6 / POPC_PER_COMPARISON_CSA / 1e9
    return min(alu, xu)


def run_ours(args, rank, local_rank, world):
    import ctypes as C

    import torch
    import torch.distributed as dist

    import racing_slam_b200 as pkg

    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: libslamgpu has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        # NCCL's version banner goes to stdout by default; stdout carries the one JSON line only
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        import datetime
        # a rank that dies must not leave the others waiting for the default 10 minutes
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank),
                                timeout=datetime.timedelta(seconds=180))

    def barrier():
        if world > 1:
            dist.barrier()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    peaks = load_peaks()
    ctx = pkg.Context(local_rank)
    lib = ctx._lib
    stream = torch.cuda.ExternalStream(ctx.stream)
    db, off, pairs = make_workload(args, rank, world)
    replicate_ms = None
    if world > 1:
        # the one exchange of the path: rank 0's key-frame database to every GPU, one NCCL broadcast
        # of the device tensor over NVLink / NVSwitch, before the clock starts.  The other ranks drop their own
        # copy and match against what arrived.
        # Every rank knows the database's shape (frames x descriptors x 32 bytes), so the receive buffer is pre-sized
        # and the exchange is ONE ncclBroadcast, timed with CUDA events on the stream it runs on.
        d_db = torch.from_numpy(db).cuda() if rank == 0 else torch.empty(db.shape, dtype=torch.uint8, device="cuda")
        warm = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda")
        for _ in range(3):
            dist.broadcast(warm, src=0)                                    # warm the communicator and its channels
        torch.cuda.synchronize()
        barrier()
        b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        b0.record()
        dist.broadcast(d_db, src=0)
        b1.record()
        torch.cuda.synchronize()
        replicate_ms = max_over_ranks(float(b0.elapsed_time(b1)))
        # The same replication as scatter + all-gather (the large-message form of a broadcast: rank 0 sends every rank
        # one 1/N piece, then all ranks exchange their pieces, so every NVLink port carries data at once instead of
        # rank 0's egress feeding a tree): timed the same way, into a second buffer, and compared with the first.
        replicate2_ms, replicate2_ok = None, None
        try:
            nbytes = int(d_db.numel())
            chunk = -(-nbytes // world)
            chunk = -(-chunk // 512) * 512
            flat = torch.empty(world * chunk, dtype=torch.uint8, device="cuda")
            mine = flat[rank * chunk:(rank + 1) * chunk]
            if rank == 0:
                flat[:nbytes].copy_(d_db.view(-1))
            pieces = [flat[r * chunk:(r + 1) * chunk] for r in range(world)] if rank == 0 else None
            w_out = torch.zeros(world * 4096, dtype=torch.uint8, device="cuda")
            w_pieces = [w_out[r * 4096:(r + 1) * 4096] for r in range(world)] if rank == 0 else None
            for _ in range(3):                                                 # warm both collectives
                dist.scatter(w_out[rank * 4096:(rank + 1) * 4096], w_pieces, src=0)
                dist.all_gather_into_tensor(w_out, w_out[rank * 4096:(rank + 1) * 4096])
            torch.cuda.synchronize()
            barrier()
            c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            c0.record()
            dist.scatter(mine, pieces, src=0)
            dist.all_gather_into_tensor(flat, mine)
            c1.record()
            torch.cuda.synchronize()
            replicate2_ms = max_over_ranks(float(c0.elapsed_time(c1)))
            replicate2_ok = bool(sum_over_ranks(float(torch.equal(flat[:nbytes], d_db.view(-1)))) == world)
            del flat, w_out
        except Exception as e:  # noqa: BLE001  (a collective this torch / NCCL build lacks: report, keep the broadcast)
            replicate2_ok = f"unavailable: {type(e).__name__}: {e}"[:200]
        if rank != 0:
            db = d_db.cpu().numpy()
        del warm
    n, K, W = args.pairs_per_step, args.steps, args.warmup
    per_pair = float(args.desc) * args.desc
    n_sm = torch.cuda.get_device_properties(local_rank).multi_processor_count

    # ---- device-resident arm: everything in HBM before the clock starts ----
    d_desc = torch.from_numpy(db.reshape(-1, 32)).cuda()
    d_off = torch.from_numpy(off).cuda()
    d_pairs = torch.from_numpy(np.stack([step_pairs(pairs, s, n) for s in range(W + K)])).cuda()
    d_out_train = torch.empty((n, args.desc), dtype=torch.int32, device="cuda")
    d_count = torch.empty(n, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()

    def step_dev(s):
        ctx.match_descriptors_batch_dev(d_desc.data_ptr(), d_off.data_ptr(), args.frames, d_pairs[s].data_ptr(), n,
                                        args.desc, MAX_DIST, RATIO, d_out_train.data_ptr(), d_count.data_ptr())

    for s in range(W):
        step_dev(s)
    ctx.sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    ctx.profile_enable(True)
    ctx.profile_read(0)
    launches0 = ctx.launches
    barrier()
    torch.cuda.synchronize()
    t_wall0 = time.time()
    e0.record(stream)
    for s in range(K):
        step_dev(W + s)
    e1.record(stream)
    torch.cuda.synchronize()
    barrier()
    t_wall1 = time.time()
    ms_dev = max_over_ranks(float(e0.elapsed_time(e1)))
    gpu_launches = ctx.launches - launches0
    scan_ms, scan_launches = ctx.profile_read(0)
    ctx.profile_enable(False)
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
    value = world * n * K * per_pair / (ms_dev * 1e-3) / 1e9
    dev_counts_last = d_count.cpu().numpy().astype(np.int64)

    # ---- e2e arm: the host-buffer C-ABI call, host<->device copies inside the timed region ----
    h_db = torch.from_numpy(db.reshape(-1, 32)).pin_memory()
    h_off = torch.from_numpy(off).pin_memory()
    h_pairs = torch.from_numpy(np.stack([step_pairs(pairs, s, n) for s in range(W + K)])).pin_memory()
    cap = n * 1024
    o_off = np.zeros(n + 1, np.int32)
    # result lists in pinned memory (what slamgpu_host_alloc hands out): the slices' copies land in them directly
    o_q = torch.zeros(cap, dtype=torch.int32).pin_memory().numpy()
    o_t = torch.zeros(cap, dtype=torch.int32).pin_memory().numpy()
    total = C.c_int64(0)
    u8p, i32p = C.POINTER(C.c_uint8), C.POINTER(C.c_int32)

    def step_host(s):
        rc = lib.slamgpu_match_descriptors_batch(
            ctx._h, C.cast(h_db.data_ptr(), u8p), C.cast(h_off.data_ptr(), i32p), args.frames,
            C.cast(h_pairs[s].data_ptr(), i32p), n, MAX_DIST, RATIO, o_off.ctypes.data_as(i32p),
            o_q.ctypes.data_as(i32p), o_t.ctypes.data_as(i32p), cap, C.byref(total))
        ctx._check(rc)
        return int(total.value)

    for s in range(W):
        step_host(s)
    barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    d2h = 0
    for s in range(K):
        d2h += 4 * (n + 1) + 8 * step_host(W + s)
    torch.cuda.synchronize()
    t_e2e = max_over_ranks(time.perf_counter() - t0)
    barrier()
    e2e_value = world * n * K * per_pair / t_e2e / 1e9
    h2d_per_step = h_db.numel() + 4 * h_off.numel() + 8 * n
    # the two API forms ran the same last step: their per-pair match counts must agree
    assert np.array_equal(np.diff(o_off.astype(np.int64)), dev_counts_last), "host and device API forms disagree"
    checksum = int(sum_over_ranks(float(dev_counts_last.sum())))
    # ... and a random sample of the last timed step's pairs against the CPU oracle (the checker, after the clock):
    # the full (query, train) lists, not just the counts
    parity_sample = None
    if rank == 0:
        from oracle import cpu_oracle
        last = step_pairs(pairs, W + K - 1, n)
        pick = np.random.default_rng(1234).choice(n, min(96, n), replace=False)
        same = True
        frames3 = db.reshape(args.frames, -1, 32)
        for p_ in pick:
            oq, ot = cpu_oracle.match_descriptors(frames3[last[p_, 0]], frames3[last[p_, 1]], MAX_DIST, RATIO)
            a, b = int(o_off[p_]), int(o_off[p_ + 1])
            same = same and np.array_equal(o_q[a:b], oq) and np.array_equal(o_t[a:b], ot)
        parity_sample = {"pairs": int(len(pick)), "identical": bool(same),
                         "against": "oracle/slam_oracle.cpp orc_match_descriptors (pinned on cv2 BFMatcher goldens), the "
                                    "(query, train) lists of randomly chosen pairs of the last timed step"}
    # ---- the same steps against a RESIDENT key-frame database (slamgpu_kfdb_*: every key frame uploaded once, when
    # it is appended, as LoopDetector keeps them; a step then sends its pair list only) ----
    KFDB = 1
    for f in range(args.frames):
        ctx.kfdb_append(KFDB, db.reshape(args.frames, -1, 32)[f])

    def step_resident(s):
        rc = lib.slamgpu_kfdb_match_pairs(ctx._h, KFDB, C.cast(h_pairs[s].data_ptr(), i32p), n, MAX_DIST, RATIO,
                                          o_off.ctypes.data_as(i32p), o_q.ctypes.data_as(i32p),
                                          o_t.ctypes.data_as(i32p), cap, C.byref(total))
        ctx._check(rc)
        return int(total.value)

    for s in range(W):
        step_resident(s)
    sampler_res = ClockSampler(local_rank) if rank == 0 else None
    barrier()
    torch.cuda.synchronize()
    ctx.trace_read()
    tw0 = time.time()
    t0 = time.perf_counter()
    d2h_res = 0
    for s in range(K):
        d2h_res += 4 * (n + 1) + 8 * step_resident(W + s)
    torch.cuda.synchronize()
    t_res_own = time.perf_counter() - t0
    tw1 = time.time()
    t_res = max_over_ranks(t_res_own)
    t_res_min = -max_over_ranks(-t_res_own)
    clocks_res = sampler_res.stop(tw0, tw1) if sampler_res else None
    trace_us, trace_calls = ctx.trace_read()
    barrier()
    assert np.array_equal(np.diff(o_off.astype(np.int64)), dev_counts_last), "resident database form disagrees"
    # one loop-closure query as LoopDetector issues it: the newest key frame against 5 retrieved candidates
    q_pairs = np.ascontiguousarray(np.stack([np.full(5, args.frames - 1), np.arange(5) * (args.frames // 7)], 1), np.int32)
    q_off = np.zeros(6, np.int32)

    def query5():
        ctx._check(lib.slamgpu_kfdb_match_pairs(ctx._h, KFDB, q_pairs.ctypes.data_as(i32p), 5, MAX_DIST, RATIO,
                                                q_off.ctypes.data_as(i32p), o_q.ctypes.data_as(i32p),
                                                o_t.ctypes.data_as(i32p), cap, C.byref(total)))
    for _ in range(5):
        query5()
    t0 = time.perf_counter()
    for _ in range(50):
        query5()
    loop_query_us = (time.perf_counter() - t0) / 50 * 1e6
    whole_job = c5_whole_job(args, ctx, pkg, lib, KFDB, rank, world, barrier, max_over_ranks, sum_over_ranks,
                             o_off, o_q, o_t, cap, n)
    ctx.kfdb_release(KFDB)
    e2e_resident = {"value": world * n * K * per_pair / t_res / 1e9, "unit": UNIT, "h2d_bytes_per_step": 8 * n,
                    "d2h_bytes_per_step": d2h_res // K, "ms_per_step": t_res / K * 1e3,
                    "api": "slamgpu_kfdb_match_pairs (key frames appended once before the clock; pair list up, match "
                           "lists down every step)",
                    "loop_query_5_candidates_us": loop_query_us,
                    "ms_per_step_fastest_rank": t_res_min / K * 1e3, "clocks_rank0": clocks_res}
    if os.environ.get("SLAMGPU_TRACE_BATCH") == "1" and trace_calls:
        # where the host time of a step goes (rank 0's own clock; SLAMGPU_TRACE_BATCH=1)
        e2e_resident["host_trace_ms_per_step_rank0"] = {
            "queueing_the_call": trace_us[0] / trace_calls / 1e3, "waiting_for_slices": trace_us[1] / trace_calls / 1e3,
            "tail_after_last_slice": trace_us[2] / trace_calls / 1e3, "whole_call": trace_us[3] / trace_calls / 1e3,
            "python_between_calls": (t_res_own * 1e6 - trace_us[3]) / trace_calls / 1e3}
    # every rank takes part in the sharded C4 batch (collectives inside): before the non-zero ranks leave
    c4_line = c4_sharded(ctx, pkg, rank, world, barrier, max_over_ranks, sum_over_ranks)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (K1 brute-force scan), live CUDA-event durations ----
    avg_ms = scan_ms / max(scan_launches, 1)
    pairs_per_launch = n * K / max(scan_launches, 1)
    cmp_per_launch = pairs_per_launch * per_pair
    achieved = cmp_per_launch / (avg_ms * 1e-3) / 1e9
    bytes_per_launch = pairs_per_launch * (2 * args.desc * 32 + args.desc * 8)
    hbm_achieved = bytes_per_launch / (avg_ms * 1e-3) / 1e9
    hbm_view = {"bound": "hbm", "achieved": hbm_achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "frac": hbm_achieved / peaks["hbm_gbs"], "peak_source": peaks["source"],
                "algorithmic_bytes_per_launch": bytes_per_launch}
    int_peak = int_peak_gcmp(n_sm, peaks["sm_max_mhz"])
    popc8_peak = popc_peak_gcmp(n_sm, peaks["sm_max_mhz"])
    if os.environ.get("SLAMGPU_BF_ENGINE") == "int":
        roofline = {
            "bound": "int-alu", "achieved": achieved, "peak": int_peak, "unit": "Gcmp/s", "frac": achieved / int_peak,
            "traffic": NCU_SCAN_DRAM_BYTES_512_PAIRS * pairs_per_launch / 512.0,
            "traffic_basis": "ncu --set full dram__bytes_read+write of one 512-pair bf_scan_kernel launch "
                             "(profiles/r01_ncu_full_prof_bf_scan_v2.txt: 7.2 MB read, 0 written), scaled to this launch's pair count",
            "kernel": "bf_scan_kernel", "launches": scan_launches, "avg_launch_ms": avg_ms,
            "share_of_step": scan_ms / ms_dev,
            "peak_basis": f"integer pipes: {n_sm} SMs x {ALU_PER_CLK_PER_SM} ALU lanes/clk/SM x {peaks['sm_max_mhz']:.0f} MHz / "
                          f"{ALU_PER_COMPARISON} ALU ops per 256-bit comparison",
            "popc8": {"peak": popc8_peak, "frac": achieved / popc8_peak}, "hbm": hbm_view,
        }
    else:
        # tensor-core engine (default for this workload): one comparison = 256 multiply-adds = 512 flop on the
        # block-scaled FP4 tensor cores.  Peak = 4 x the measured dense bf16 rate of MEASURED_PEAKS.json (fp4 : bf16
        # = 9 : 2.25 PFLOP/s nominal); burst figure, the conservative denominator.
        tflops = achieved * 1e9 * 512.0 / 1e12
        peak_tf = 4.0 * peaks["bf16_tflops"]
        roofline = {
            "bound": "tensor", "achieved": tflops, "peak": peak_tf, "unit": "TFLOP/s", "frac": tflops / peak_tf,
            "traffic": NCU_TC4_DRAM_BYTES_1024_PAIRS * pairs_per_launch / 1024.0,
            "traffic_basis": "ncu --set full dram__bytes_read+write of one 1024-pair bf_tc4_scan_kernel launch "
                             "(profiles/r02_final_ncu_full_bf_tc4_scan.txt: 27.5 MB read + 2.5 MB written), scaled to this launch's pair count",
            "kernel": "bf_tc4_scan_kernel", "launches": scan_launches, "avg_launch_ms": avg_ms,
            "share_of_step": scan_ms / ms_dev, "achieved_gcmp_s": achieved,
            "peak_basis": f"4 x measured dense bf16 ({peaks['bf16_tflops']:.0f} TFLOP/s burst, {peaks['source']}); algorithmic work "
                          "= 512 flop per 256-bit comparison.  The kernel also issues one constant MMA per 8 useful ones and pads "
                          "4000 train rows to 9 tiles of 480, so its tensor-pipe time is 1.215x the algorithmic one (MMA-only floor of this "
                          "instruction mix: 0.89 of the measured launch time, DESIGN.md section 4 K7b)",
            "vs_sustained_bf16_x4": tflops / (4.0 * peaks["bf16_tflops_sustained"]),
            "vs_integer_pipe_roofline": {"note": "the XOR+POPC formulation's ceilings (SURVEY 8d), for scale",
                                         "int_alu_peak_gcmp_s": int_peak, "x_int_alu": achieved / int_peak,
                                         "popc8_peak_gcmp_s": popc8_peak, "x_popc8": achieved / popc8_peak},
            "hbm": hbm_view,
        }

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_dev / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8" if os.environ.get("SLAMGPU_BF_ENGINE") == "int" else "e2m1 (fp4 products of +-1, exact fp32 accumulation)",
        "data": "synthetic",
        "config": config_of(args),
        "step": {"pairs_per_step": n, "pairs_per_step_per_gpu": n, "comparisons_per_step": world * n * per_pair,
                 "sharding": f"frame pairs block-cyclic over {world} rank(s), database replicated, no collective",
                 "l2": f"database {db.nbytes / 2**20:.0f} MiB (>= 126 MB L2), different pairs every step"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d_per_step),
                "d2h_bytes_per_step": int(d2h // K), "ms_per_step": t_e2e / K * 1e3,
                "api": "slamgpu_match_descriptors_batch (host buffers; descriptors re-uploaded every step)"},
        "e2e_resident_database": e2e_resident,
        "gpu_launches": int(gpu_launches), "roofline": roofline, "clocks": clocks,
        "matches_last_step_all_ranks": checksum, "parity_sample": parity_sample,
    }
    line["c4_local_window_batch"] = c4_line
    line["c5_whole_job_strong"] = whole_job
    if replicate_ms is not None:
        line["replicate"] = {"what": "key-frame database resident on GPU 0 -> all ranks, one pre-sized ncclBroadcast of the "
                                     "device tensor (CUDA events on the collective's stream, max over ranks; outside the "
                                     "timed region)", "bytes": int(db.nbytes), "ms": replicate_ms,
                             "gb_s": db.nbytes / (replicate_ms * 1e-3) / 1e9,
                             "scatter_allgather": {"what": "the same replication as ncclScatter-by-send/recv of 1/N pieces + one "
                                                           "ncclAllGather (in place), same clock",
                                                   "ms": replicate2_ms,
                                                   "gb_s": (db.nbytes / (replicate2_ms * 1e-3) / 1e9) if replicate2_ms else None,
                                                   "identical_to_broadcast": replicate2_ok}}
    if world == 1:
        line["cpu_baseline"] = cpu_baseline(args, db)
        if not args.no_also:
            line["also"] = also_measurements(ctx, pkg, torch, stream, n_sm, peaks)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def c5_whole_job(args, ctx, pkg, lib, kfdb, rank, world, barrier, max_over_ranks, sum_over_ranks, o_off, o_q, o_t, cap, n):
    """BASELINE configs[4] as ONE job, strong-scaled: every ordered pair (query > candidate) of the key-frame database
    -- 499,500 pairs = 7.99e12 comparisons at the full size -- sharded block-cyclically over the ranks, against the
    resident database, match lists returned to pinned host arrays.  Host clock between barriers, max over ranks; the
    total match count is a checksum that must not depend on the number of ranks or on the slicing."""
    import ctypes as C
    i32p = C.POINTER(C.c_int32)
    all_p = pkg.sharding.all_pairs(args.frames)
    mine = pkg.sharding.shard_pairs(all_p, rank, world, block=64)
    total = C.c_int64(0)
    off = np.zeros(n + 1, np.int32)
    times, matches = [], 0
    for rep in range(2):                            # the second pass is the one reported (first: warm allocations)
        barrier()
        t0 = time.perf_counter()
        matches = 0
        for s0 in range(0, len(mine), n):
            chunk = np.ascontiguousarray(mine[s0:s0 + n])
            ctx._check(lib.slamgpu_kfdb_match_pairs(ctx._h, kfdb, chunk.ctypes.data_as(i32p), len(chunk), MAX_DIST, RATIO,
                                                    off.ctypes.data_as(i32p), o_q.ctypes.data_as(i32p),
                                                    o_t.ctypes.data_as(i32p), cap, C.byref(total)))
            matches += int(total.value)
        times.append(max_over_ranks(time.perf_counter() - t0))
    secs = times[-1]
    cmp_total = float(len(all_p)) * args.desc * args.desc
    return {"pairs": int(len(all_p)), "pairs_this_rank": int(len(mine)), "seconds": secs, "gcmp_s": cmp_total / secs / 1e9,
            "matches_total": int(sum_over_ranks(float(matches))), "scaling": "strong",
            "api": "slamgpu_kfdb_match_pairs in calls of %d pairs (resident database; pair lists up, match lists down "
                   "into pinned arrays)" % n,
            "sharding": f"pair list block-cyclic (blocks of 64) over {world} rank(s), no collective",
            "clock": "host wall between barriers, max over ranks"}


def c4_sharded(ctx, pkg, rank, world, barrier, max_over_ranks, sum_over_ranks):
    """BASELINE configs[3]: the fuse window -- 20 key frames x 50k map points on a frozen
    snapshot, replace = true -- sharded by key frame (round-robin, sharding.shard_keyframes), point set
    replicated (same seed on every rank), no collective on the data path.  'strong' = the 20 key frames of the
    config split over the ranks; 'weak' = 20 key frames per rank.  Host-buffer call (poses up, matches down),
    host clock between barriers, max over ranks, median of the repetitions."""
    syn = pkg.synthetic
    cam = syn.KITTI
    sc = syn.projection_scene(50000, 2000, 4)
    prm = pkg.ProjParams(cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["width"], cam["height"], 20.0, MAX_DIST, 0.5,
                         2.0, 1.25, 1)
    ctx.points_upload(940, sc.pos, sc.normal, sc.range, sc.obs_offset, sc.obs_desc)
    out = {}
    for mode, n_kf in (("strong", 20), ("weak", 20 * world)):
        rng = np.random.default_rng(4)
        xys = [syn.keypoints_integer(2000, cam["width"], cam["height"], 700 + k) for k in range(n_kf)]
        descs = [sc.kp_desc[rng.permutation(2000)] for _ in range(n_kf)]
        poses_all = np.stack([syn.pose_colmajor(syn.make_pose(rng)) for _ in range(n_kf)])
        mine = pkg.sharding.shard_keyframes(n_kf, rank, world)
        ids = []
        for j, k in enumerate(mine):
            ctx.frame_upload(2000 + j, xys[k], descs[k], None, None)
            ids.append(2000 + j)
        poses = poses_all[mine]
        nk = [2000] * len(ids)
        ts, matches = [], 0
        for rep in range(13):
            barrier()
            t0 = time.perf_counter()
            res = ctx.match_projection_batch(ids, nk, 940, poses, prm) if ids else []
            dt = max_over_ranks(time.perf_counter() - t0)
            if rep >= 3:
                ts.append(dt)
            matches = sum(int((p >= 0).sum()) for p, _ in res)
        s = float(np.median(ts))
        out[mode] = {"key_frames_total": n_kf, "key_frames_this_rank": len(ids), "ms_per_batch": s * 1e3,
                     "keyframes_per_s": n_kf / s, "point_projections_per_s": n_kf * 50000 / s,
                     "matches_all_ranks": int(sum_over_ranks(float(matches)))}
        for i in ids:
            ctx.frame_release(i)
    ctx.points_release(940)
    out["sharding"] = f"key frames round-robin over {world} rank(s), 50k-point set replicated, no collective; copies inside"
    return out


def also_measurements(ctx, pkg, torch, stream, n_sm, peaks):
    """Secondary numbers on the other BASELINE configs (N=1 only; not the headline):
    C1 single 2000x2000 pair, 8k x 8k device-resident kNN-2 (the north-star roofline bar),
    C2 projection-guided search per frame, C4 key-frame batch.  All through the C ABI."""
    syn = pkg.synthetic
    out = {}

    def wall(fn, reps, warm=3):
        for _ in range(warm):
            fn()
        ts = []
        for _ in range(reps):
            t = time.perf_counter()
            fn()
            ts.append(time.perf_counter() - t)
        return float(np.median(ts))

    # C1: host-buffer calls, H2D + D2H inside
    q, t = syn.bf_correlated(2000, 2000, 1)
    s = wall(lambda: ctx.match_descriptors(q, t, MAX_DIST, RATIO), 50)
    out["c1_match_descriptors_2000x2000"] = {"us_per_call": s * 1e6, "gcmp_s": 4e6 / s / 1e9, "copies": "inside"}
    s = wall(lambda: ctx.match_features(t, q, MAX_DIST), 50)
    out["c1_match_features_2000x2000"] = {"us_per_call": s * 1e6, "gcmp_s": 4e6 / s / 1e9, "copies": "inside",
                                         "note": "cross-check: both directions are needed"}

    # 8k x 8k kNN-2, device-resident, CUDA events; L2 flushed between iterations
    rng = np.random.default_rng(8)
    dq = torch.from_numpy(rng.integers(0, 256, (8000, 32), dtype=np.uint8)).cuda()
    dt = torch.from_numpy(rng.integers(0, 256, (8000, 32), dtype=np.uint8)).cuda()
    idx = torch.empty((8000, 2), dtype=torch.int32, device="cuda")
    dist_ = torch.empty((8000, 2), dtype=torch.int32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    for _ in range(3):
        ctx.knn2_dev(dq.data_ptr(), 8000, dt.data_ptr(), 8000, idx.data_ptr(), dist_.data_ptr())
    ctx.sync()
    ctx.profile_enable(True)
    ctx.profile_read(0)
    tot = 0.0
    for _ in range(20):
        flush.fill_(1)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        ctx.knn2_dev(dq.data_ptr(), 8000, dt.data_ptr(), 8000, idx.data_ptr(), dist_.data_ptr())
        b.record(stream)
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    scan_ms, scan_n = ctx.profile_read(0)
    ctx.profile_enable(False)
    g_call = 64e6 / (tot / 20 * 1e-3) / 1e9
    g_kernel = 64e6 / (scan_ms / max(scan_n, 1) * 1e-3) / 1e9
    peak = popc_peak_gcmp(n_sm, peaks["sm_max_mhz"])
    ipeak = int_peak_gcmp(n_sm, peaks["sm_max_mhz"])
    out["knn2_8000x8000_device"] = {"us_per_call": tot / 20 * 1e3, "gcmp_s": g_call, "scan_kernel_gcmp_s": g_kernel,
                                   "popc8_roofline_frac_call": g_call / peak, "popc8_roofline_frac_kernel": g_kernel / peak,
                                   "int_roofline_frac_kernel": g_kernel / ipeak,
                                   "l2": "flushed (256 MiB fill) before every call"}
    # C3 frame-to-frame: 8000 x 8000 cross-check (match_features), host buffers, copies inside
    prev = rng.integers(0, 256, (8000, 32), dtype=np.uint8)
    cur = syn.next_frame_descriptors(prev, 33)
    s = wall(lambda: ctx.match_features(prev, cur, MAX_DIST), 30)
    out["c3_match_features_8000x8000"] = {"us_per_call": s * 1e6, "frames_per_s": 1.0 / s, "gcmp_s": 64e6 / s / 1e9,
                                          "copies": "inside", "note": "one fused scan: row top-1 + column minima"}
    ci = torch.empty(8000, dtype=torch.int32, device="cuda")
    cd = torch.empty(8000, dtype=torch.int32, device="cuda")
    tot = 0.0
    for it in range(23):
        flush.fill_(1)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        ctx.crosscheck_dev(dq.data_ptr(), 8000, dt.data_ptr(), 8000, ci.data_ptr(), cd.data_ptr())
        b.record(stream)
        torch.cuda.synchronize()
        if it >= 3:
            tot += a.elapsed_time(b)
    out["crosscheck_8000x8000_device"] = {"us_per_call": tot / 20 * 1e3, "gcmp_s": 64e6 / (tot / 20 * 1e-3) / 1e9}
    del flush

    # C2: resident frame + point set; per frame: pose in, proposals out
    cam = syn.KITTI
    sc = syn.projection_scene(10000, 2000, 2)
    ctx.frame_upload(900, sc.kp_xy, sc.kp_desc, None, sc.kp_matched)
    ctx.points_upload(900, sc.pos, sc.normal, sc.range, sc.obs_offset, sc.obs_desc)
    for radius in (15.0, 20.0):
        prm = pkg.ProjParams(cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["width"], cam["height"], radius,
                             MAX_DIST, 0.5, 2.0, 1.25, 0)
        s = wall(lambda: ctx.match_projection(900, 900, 2000, sc.pose, prm, sc.skip, None), 200)
        ctx.profile_enable(True)
        ctx.profile_read(1)
        for _ in range(50):
            ctx.match_projection(900, 900, 2000, sc.pose, prm, sc.skip, None)
        k_ms, k_n = ctx.profile_read(1)
        ctx.profile_enable(False)
        nbytes = (len(sc.pos) * 37 + len(sc.obs_desc) * 32 + 2000 * 45 + 2000 * 8)
        out[f"c2_projection_10k_points_2000_kp_r{radius:g}"] = {
            "frames_per_s": 1.0 / s, "us_per_frame": s * 1e6, "match_kernel_us": k_ms / max(k_n, 1) * 1e3,
            "algorithmic_bytes": nbytes, "kernel_gb_s": nbytes / (k_ms / max(k_n, 1) * 1e-3) / 1e9,
            "hbm_frac": nbytes / (k_ms / max(k_n, 1) * 1e-3) / 1e9 / peaks["hbm_gbs"], "copies": "inside"}
    ctx.frame_release(900)
    ctx.points_release(900)

    # C4: 20 key frames x 50k points, frozen snapshot, replace = true
    sc = syn.projection_scene(50000, 2000, 4)
    rng = np.random.default_rng(4)
    ids, nkps, poses = [], [], []
    for k in range(20):
        xy = syn.keypoints_integer(2000, cam["width"], cam["height"], 700 + k)
        ctx.frame_upload(1000 + k, xy, sc.kp_desc[rng.permutation(2000)], None, None)
        ids.append(1000 + k), nkps.append(2000), poses.append(syn.pose_colmajor(syn.make_pose(rng)))
    ctx.points_upload(901, sc.pos, sc.normal, sc.range, sc.obs_offset, sc.obs_desc)
    prm = pkg.ProjParams(cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["width"], cam["height"], 20.0, MAX_DIST, 0.5,
                         2.0, 1.25, 1)
    poses = np.stack(poses)
    s = wall(lambda: ctx.match_projection_batch(ids, nkps, 901, poses, prm), 30)
    out["c4_fuse_batch_20_keyframes_x_50k_points"] = {"ms_per_batch": s * 1e3, "keyframes_per_s": 20.0 / s,
                                                      "point_projections_per_s": 20 * 50000 / s, "copies": "inside"}
    for k in ids:
        ctx.frame_release(k)
    ctx.points_release(901)

    # C3: the tracking sequence, shortened 10x in frames with 10x the points per frame so that the map
    # still grows 0 -> 200k (tools/bench_c3.py runs the full 2000 frames; profiles/ holds its output)
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import bench_c3
    bench_c3.run_sequence(ctx, pkg, 10, 8000, 1000, 3, 0)
    out["c3_tracking_sequence_200_frames"] = bench_c3.run_sequence(ctx, pkg, 200, 8000, 1000, 3, 3, c3_cpu_arm)
    out["k7b_hit_density_sweep"] = hit_density_sweep(ctx, pkg, torch, stream, peaks)
    out["adapter_track_frame"] = adapter_track_frame(pkg)
    out["c_abi_call_latency"] = c_abi_call_latency(pkg)
    out["n4_bag_of_words"] = bow_measurements(ctx, pkg, torch, stream, peaks)
    out["n3_pnp_scoring"] = pnp_measurements(ctx, pkg, torch, stream)
    return out


def hit_density_sweep(ctx, pkg, torch, stream, peaks):
    """The FP4 filtered scan (K7b) decodes accumulators only where a threshold bit is set, so its speed depends on how
    many (query, train) rows fall within the distance cut.  64 key frames x 4000 descriptors, all 2016 ordered pairs
    per launch, device-resident; the share of descriptors drawn from the common landmark pool sets the hit density."""
    out = []
    frames, desc = 64, 4000
    pairs = pkg.sharding.all_pairs(frames)
    off = (np.arange(frames + 1, dtype=np.int64) * desc).astype(np.int32)
    d_off, d_pairs = torch.from_numpy(off).cuda(), torch.from_numpy(pairs).cuda()
    d_out = torch.empty((len(pairs), desc), dtype=torch.int32, device="cuda")
    d_cnt = torch.empty(len(pairs), dtype=torch.int32, device="cuda")
    for shared in (0.0, 0.03, 0.1, 0.3, 0.55, 1.0):
        db = pkg.synthetic.keyframe_db_fast(frames, desc, 77, shared)
        d_db = torch.from_numpy(db.reshape(-1, 32)).cuda()
        run = lambda: ctx.match_descriptors_batch_dev(d_db.data_ptr(), d_off.data_ptr(), frames, d_pairs.data_ptr(),  # noqa: E731
                                                      len(pairs), desc, MAX_DIST, RATIO, d_out.data_ptr(), d_cnt.data_ptr())
        for _ in range(3):
            run()
        ctx.sync()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(stream)
        for _ in range(10):
            run()
        b.record(stream)
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 10
        g = len(pairs) * float(desc) * desc / (ms * 1e-3) / 1e9
        out.append({"shared_descriptor_share": shared, "query_rows_matched_share": float(d_cnt.sum().item()) / (len(pairs) * desc),
                    "ms_per_launch_chain": ms, "gcmp_s": g, "tensor_frac_of_4x_bf16_burst": g * 512e9 / 1e12 / (4 * peaks["bf16_tflops"])})
    return out


def c_abi_call_latency(pkg, keep_inputs=False):
    """The single-call paths as a C++ host sees them (tools/call_latency.cpp: plain g++ against include/slamgpu.h,
    steady_clock around each call, copies inside, no interpreter): the same synthetic scenes as the Python legs
    above, whose figures include numpy / ctypes marshalling."""
    import struct
    tool = os.path.join(ROOT, "tools", "call_latency")
    if not os.path.exists(tool):
        return {"unavailable": "tools/call_latency is not built (__graft_entry__.build() makes it)"}
    syn = pkg.synthetic
    blobs = {}
    q, t = syn.bf_correlated(2000, 2000, 1)
    blobs["c1_q"], blobs["c1_t"] = q, t
    rng = np.random.default_rng(5)
    prev = rng.integers(0, 256, (8000, 32), dtype=np.uint8)
    blobs["c3_prev"], blobs["c3_cur"] = prev, syn.next_frame_descriptors(prev, 33)
    cam = syn.KITTI
    sc = syn.projection_scene(10000, 2000, 2)
    prm = pkg.ProjParams(cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["width"], cam["height"], 20.0, MAX_DIST, 0.5,
                         2.0, 1.25, 0)
    blobs.update(c2_kp_xy=sc.kp_xy.astype(np.float32), c2_kp_desc=sc.kp_desc, c2_kp_matched=sc.kp_matched.astype(np.uint8),
                 c2_pos=sc.pos.astype(np.float32), c2_normal=sc.normal.astype(np.float32), c2_range=sc.range.astype(np.float32),
                 c2_obs_offset=sc.obs_offset.astype(np.int32), c2_obs_desc=sc.obs_desc,
                 c2_pose=np.ascontiguousarray(sc.pose, np.float32).reshape(16), c2_skip=sc.skip.astype(np.uint8),
                 c2_params=np.frombuffer(bytes(prm), np.uint8))
    blobs["db_desc"] = syn.keyframe_db_fast(66, 4000, 77, 0.3)
    blobs["db_rows"] = np.array([4000], np.int32)
    path = os.path.join(ROOT, "gpurun_out", "call_latency_inputs.blob")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "wb") as f:
        for name, arr in blobs.items():
            raw = np.ascontiguousarray(arr).tobytes()
            f.write(struct.pack("<I", len(name)) + name.encode() + struct.pack("<Q", len(raw)) + raw)
    if keep_inputs:        # for a profiler run of the tool itself (tools/run_call_latency.py --keep)
        return {"inputs": path}
    r = subprocess.run([tool, path, "300"], capture_output=True, text=True)
    os.remove(path)
    if r.returncode != 0:
        return {"unavailable": f"call_latency failed: {r.stderr[-300:]}"}
    res = json.loads(r.stdout)
    res["clock"] = "std::chrono::steady_clock around each C-ABI call in a C++ host, median of 300, copies inside"
    return res


def adapter_track_frame(pkg):
    """One tracked frame THROUGH THE C++ ADAPTER (racing-slam_b200/adapter/MapMatcherGpu, compiled against the mock of
    the reference's data model): MapMatcherGpu::track on a 200k-point map with a mutation mix (a key-frame insertion
    every 5th frame: new points, new observations, culled points, bundle-adjusted positions and poses), wall clock,
    the resident mirror (deltas) against per-call uploads of the whole map; beside it the oracle's two projection calls
    on a scene of the same size (1 core, like the reference)."""
    from oracle import cpu_oracle
    driver = os.path.join(ROOT, "tests", "adapter_mock", "adapter_driver")
    out_path = os.path.join(ROOT, "gpurun_out", "adapter_bench.json")
    os.makedirs(os.path.dirname(out_path), exist_ok=True)
    if not os.path.exists(driver):
        return {"unavailable": "tests/adapter_mock/adapter_driver is not built (__graft_entry__.build() makes it)"}
    r = subprocess.run([driver, "--bench", "200000", "8000", "60", out_path], capture_output=True, text=True)
    if r.returncode != 0:
        return {"unavailable": f"adapter driver failed: {r.stderr[-300:]}"}
    with open(out_path) as f:
        res = json.load(f)
    syn = pkg.synthetic
    cam = syn.HD1080
    sc = syn.projection_scene(200000, 8000, 3, cam=cam)
    oprm = cpu_oracle.ProjParams(cam["fx"], cam["fy"], cam["cx"], cam["cy"], cam["width"], cam["height"], 20.0, MAX_DIST,
                                 0.5, 2.0, 1.25, 0)
    perm = cpu_oracle.kd_build(sc.kp_xy)
    in_kf = np.zeros(200000, np.uint8)
    in_kf[::66] = 1
    t0 = time.perf_counter()
    for skip in ((1 - in_kf).astype(np.uint8), np.zeros(200000, np.uint8)):
        cpu_oracle.match_projection(sc.kp_xy, sc.kp_desc, perm, sc.kp_matched, sc.pos, sc.normal, sc.range, sc.obs_offset,
                                    sc.obs_desc, skip, None, sc.pose, oprm, want_stats=False)
    res["oracle_projection_pair_ms_1_core"] = (time.perf_counter() - t0) * 1e3
    res["note"] = ("track_ms_* exclude the add_map_match loops, which are the reference's own host code "
                   "(O(matches x keypoints), add_map_match_loops_ms) whichever matcher produced the lists")
    return res


def pnp_measurements(ctx, pkg, torch, stream):
    """the scoring step of verify_pnp (src/LoopDetector.cpp:203-214: 200 RANSAC models, reprojection
    error 4.0).  One problem through the host-buffer call (copies inside) and a 4000-problem batch device-resident;
    the CPU port (oracle, all cores over models) beside them.  An evaluation = one correspondence under one model:
    23 FP64 multiplies / adds, one FP64 division, five conversions, five FP32 operations."""
    from oracle import cpu_oracle
    syn = pkg.synthetic
    K4 = syn.PNP_K4
    out = {}
    cores = os.cpu_count()

    def wall(fn, reps, warm=3):
        for _ in range(warm):
            fn()
        ts = []
        for _ in range(reps):
            t = time.perf_counter()
            fn()
            ts.append(time.perf_counter() - t)
        return float(np.median(ts))

    objs, imgs, models = syn.pnp_problems(1, [300], [200])
    s = wall(lambda: ctx.pnp_ransac_score(objs, imgs, models, K4, 4.0), 50)
    got = ctx.pnp_ransac_score(objs, imgs, models, K4, 4.0)[0]
    hp = np.zeros(200, np.int32)
    t_cpu = wall(lambda: cpu_oracle.pnp_score_batch(objs[0], imgs[0], [0], [300], models[0], hp, K4, 4.0, 1), 20)
    ocnt = cpu_oracle.pnp_score_batch(objs[0], imgs[0], [0], [300], models[0], hp, K4, 4.0, 1)
    out["one_problem_300x200"] = {"us_per_call": s * 1e6, "copies": "inside (two syncs: counts down, choice up, mask down)",
                                  "cpu_port_us_1_core": t_cpu * 1e6,
                                  "identical_to_oracle": bool(np.array_equal(got[0], ocnt) and
                                                              got[1] == cpu_oracle.ransac_select(ocnt, 300)[0])}

    n_prob, n_mod = 4000, 200
    rng = np.random.default_rng(9)
    sizes = [int(x) for x in rng.integers(40, 560, n_prob)]
    objs, imgs, models = syn.pnp_problems(2, sizes, [n_mod] * n_prob)
    obj, img, rt = np.concatenate(objs), np.concatenate(imgs), np.concatenate(models)
    cnt = np.array(sizes, np.int32)
    begin = (np.cumsum(cnt) - cnt).astype(np.int32)
    hp = np.repeat(np.arange(n_prob, dtype=np.int32), n_mod)
    evals = int(cnt.astype(np.int64).sum()) * n_mod
    d = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    d_obj, d_img, d_rt, d_begin, d_cnt, d_hp = d(obj), d(img), d(rt), d(begin), d(cnt), d(hp)
    d_good = torch.empty(len(hp), dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    run = lambda: ctx.pnp_score_dev(d_obj.data_ptr(), d_img.data_ptr(), d_begin.data_ptr(), d_cnt.data_ptr(), n_prob,
                                    d_rt.data_ptr(), d_hp.data_ptr(), len(hp), K4, 4.0, d_good.data_ptr())
    for _ in range(2):
        run()
    ctx.sync()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(5):
        run()
    b.record(stream)
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 5
    sample = 400 * n_mod                                   # the first 400 problems
    t = time.perf_counter()
    ocnt = cpu_oracle.pnp_score_batch(obj, img, begin, cnt, rt[:sample], hp[:sample], K4, 4.0, cores)
    t_cpu = time.perf_counter() - t
    sample_evals = int(cnt[:400].astype(np.int64).sum()) * n_mod
    out["batch_4000_problems_x_200_models_device"] = {
        "ms": ms, "evaluations": evals, "evaluations_per_s": evals / (ms * 1e-3),
        "fp64_ops_per_s": evals * 33 / (ms * 1e-3),
        "fp64_ops_note": "14 DMUL + 11 DADD + 8 DFMA of the IEEE division per evaluation (SASS count)",
        "algorithmic_bytes": int(len(obj) * 20 + len(rt) * 100 + len(rt) * 4),
        "cpu_port_evaluations_per_s": sample_evals / t_cpu, "cpu_port_cores": cores,
        "cpu_port_sample": "first 400 problems (80000 models)",
        "identical_to_oracle_on_sample": bool(np.array_equal(d_good[:sample].cpu().numpy(), ocnt))}
    return out


def bow_measurements(ctx, pkg, torch, stream, peaks):
    """LoopDetector's retrieval step on an ORB-vocabulary-sized tree (k = 10, L = 6, 1.1 M nodes,
    synthetic): bow_of for one 8000-feature key frame through the host-buffer call, the word lookup for a
    1000 x 4000 batch device-resident, and score_candidates against 1000 key frames; CPU legs beside them."""
    from oracle import cpu_oracle
    syn = pkg.synthetic
    out = {}
    voc = syn.random_vocabulary(10, 6, 6)
    n_nodes = len(voc["desc"])
    ctx.vocab_upload(50, voc["child_offset"], voc["child_ids"], voc["desc"], voc["weight"], voc["word_id"])
    ov = cpu_oracle.Vocabulary(voc["child_offset"], voc["child_ids"], voc["desc"], voc["weight"], voc["word_id"])
    bytes_per_desc = 32 + 6 * (16 + 32 * 10)            # descriptor + per level: one node record + ten child descriptors

    def wall(fn, reps, warm=3):
        for _ in range(warm):
            fn()
        ts = []
        for _ in range(reps):
            t = time.perf_counter()
            fn()
            ts.append(time.perf_counter() - t)
        return float(np.median(ts))

    # one key frame, host buffers, copies inside
    q = syn.vocabulary_queries(voc, 8000, 7)
    s = wall(lambda: ctx.vocab_bow(50, q), 50)
    ctx.profile_enable(True)
    ctx.profile_read(2)
    for _ in range(20):
        ctx.vocab_bow(50, q)
    k_ms, k_n = ctx.profile_read(2)
    ctx.profile_enable(False)
    got = ctx.vocab_bow(50, q)[0]
    t = time.perf_counter()
    word, w = cpu_oracle.voc_transform_each(ov, q)
    ob = cpu_oracle.voc_bow(word, w)
    t_port = time.perf_counter() - t
    out["bow_of_8000_features"] = {
        "us_per_call": s * 1e6, "transform_kernel_us": k_ms / max(k_n, 1) * 1e3, "copies": "inside",
        "vocabulary": f"k=10 L=6 synthetic, {n_nodes} nodes ({n_nodes * 32 / 1e6:.0f} MB of node descriptors)",
        "cpu_port_us": t_port * 1e6, "cpu_port_cores": os.cpu_count(),
        "identical_to_oracle": bool(np.array_equal(got[0], ob[0]) and np.array_equal(got[1], ob[1]))}
    if cpu_oracle.ref_dbow2() is not None:
        # the reference's own DBoW2 on a tree of its own (single-threaded, like bow_of in the reference)
        rng = np.random.default_rng(3)
        pool = rng.integers(0, 256, (4000, 32), dtype=np.uint8)
        imgs = [syn.flip_bits(pool[rng.choice(4000, 500, replace=False)], rng.integers(0, 40, 500), rng) for _ in range(40)]
        rv = cpu_oracle.RefVocabulary(10, 4, np.concatenate(imgs), np.arange(41) * 500)
        qq = syn.flip_bits(pool[rng.choice(4000, 8000)], rng.integers(0, 60, 8000), rng)
        t = time.perf_counter()
        rv.transform_bow(qq)
        out["bow_of_8000_features"]["reference_dbow2_us_k10_L4"] = (time.perf_counter() - t) * 1e6
        rv.close()

    # the word lookup for a C5-sized batch, device-resident (1000 key frames x 4000 descriptors)
    n_big = 4_000_000
    rng = np.random.default_rng(8)
    leaves = np.nonzero(voc["word_id"] >= 0)[0]
    base = voc["desc"][rng.choice(leaves, 200_000)]
    qbig = np.repeat(base, n_big // len(base), axis=0) ^ rng.integers(0, 256, (n_big, 32), dtype=np.uint8) & \
        rng.integers(0, 256, (n_big, 32), dtype=np.uint8) & rng.integers(0, 256, (n_big, 32), dtype=np.uint8)
    dq = torch.from_numpy(np.ascontiguousarray(qbig)).cuda()
    dword = torch.empty(n_big, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    for _ in range(2):
        ctx.vocab_transform_dev(50, dq.data_ptr(), n_big, dword.data_ptr())
    ctx.sync()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for _ in range(5):
        ctx.vocab_transform_dev(50, dq.data_ptr(), n_big, dword.data_ptr())
    b.record(stream)
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 5
    sample = 200_000
    t = time.perf_counter()
    ow, _ = cpu_oracle.voc_transform_each(ov, qbig[:sample])
    t_cpu = time.perf_counter() - t
    out["transform_4M_descriptors_device"] = {
        "ms": ms, "descriptors_per_s": n_big / (ms * 1e-3), "algorithmic_bytes_per_descriptor": bytes_per_desc,
        "gb_s_algorithmic": n_big * bytes_per_desc / (ms * 1e-3) / 1e9,
        "x_hbm_peak": n_big * bytes_per_desc / (ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
        "note": "the tree (49 MB with its node records) stays in the 126 MB L2, so the algorithmic bytes are served by L2 "
                "and may exceed the HBM figure; ncu: issue slots 56 %, L1/TEX 51 % busy (profiles/r01_ncu_full_voc_transform_1M.txt)",
        "cpu_port_descriptors_per_s": sample / t_cpu, "cpu_port_cores": os.cpu_count(),
        "identical_to_oracle_on_sample": bool(np.array_equal(dword[:sample].cpu().numpy(), ow))}
    del dq, dword

    # score_candidates: one key frame against 1000 earlier ones
    per = 4000
    qdb = qbig[:1000 * per]
    off = (np.arange(1001) * per).astype(np.int32)
    db = ctx.vocab_bow(50, qdb, off)
    query = ctx.vocab_bow(50, np.concatenate([qdb[:per // 2], qdb[7 * per:7 * per + per // 2]]))[0]
    s = wall(lambda: ctx.vocab_score_l1(query, db), 10)
    t = time.perf_counter()
    want = np.array([cpu_oracle.voc_score_l1(*query, *d) for d in db])
    t_cpu = time.perf_counter() - t
    for vec in db:
        ctx.bowdb_append(51, vec)
    s_res = wall(lambda: ctx.bowdb_score_l1(51, query), 50)
    out["score_l1_1000_keyframes"] = {"us_per_call_resident_database": s_res * 1e6,
                                      "us_per_call_database_uploaded_per_call": s * 1e6,
                                      "copies": "inside (query up, 1000 scores down)",
                                      "cpu_port_us": t_cpu * 1e6, "cpu_port_note": "one core, called per vector from Python",
                                      "identical_to_oracle": bool(np.array_equal(ctx.bowdb_score_l1(51, query), want))}
    ctx.bowdb_release(51)
    ctx.vocab_release(50)
    return out


def c3_cpu_arm(pkg, seq, kept, prm):
    """The same four steps on the host for the sampled frames; also the parity check of those frames."""
    from oracle import cpu_oracle
    try:
        import cv2
        cv2.setNumThreads(os.cpu_count() or 1)
    except Exception:           # the GPU box may not have cv2: the oracle port stands in
        cv2 = None
    oprm = cpu_oracle.ProjParams(prm.fx, prm.fy, prm.cx, prm.cy, prm.width, prm.height, prm.radius, prm.max_dist,
                                 prm.min_cos, prm.nearer, prm.further, 0)
    p = seq.pool
    t_bf = t_kd = t_proj = 0.0
    parity_ok = True
    for f, (prev, cur, (ft, fq), res) in sorted(kept.items()):
        n = seq.map_size(f)
        t0 = time.perf_counter()
        if cv2 is not None:
            ms = cv2.BFMatcher(cv2.NORM_HAMMING, True).knnMatch(cur["kp_desc"], prev["kp_desc"], 1)   # FeatureExtractor.cpp:20-21
            pairs = [(m[0].trainIdx, m[0].queryIdx) for m in ms if m and m[0].distance < MAX_DIST]
            o_t = np.array([a for a, _ in pairs], np.int32)
            o_q = np.array([b for _, b in pairs], np.int32)
        else:
            o_t, o_q = cpu_oracle.match_features(prev["kp_desc"], cur["kp_desc"], MAX_DIST)
        t1 = time.perf_counter()
        perm = cpu_oracle.kd_build(cur["kp_xy"])
        t2 = time.perf_counter()
        pt = np.zeros(n, np.uint8)
        pt[cur["matched_points"]] = 1
        in_kf = np.zeros(n, np.uint8)
        in_kf[cur["kf_points"]] = 1
        kpm = cur["kp_matched"].copy()
        off = p.obs_offset[:n + 1]
        args = (p.pos[:n], p.normal[:n], p.range[:n], off, p.obs_desc[:int(off[-1])])
        r1 = cpu_oracle.match_projection(cur["kp_xy"], cur["kp_desc"], perm, kpm, *args, (pt | (1 - in_kf)).astype(np.uint8),
                                         None, cur["pose"], oprm, want_stats=False)
        hit = r1.point >= 0
        kpm[hit] = 1
        pt[r1.point[hit]] = 1
        r2 = cpu_oracle.match_projection(cur["kp_xy"], cur["kp_desc"], perm, kpm, *args, pt, None, cur["pose"], oprm,
                                         want_stats=False)
        t3 = time.perf_counter()
        t_bf += t1 - t0
        t_kd += t2 - t1
        t_proj += t3 - t2
        parity_ok &= bool(np.array_equal(ft, o_t) and np.array_equal(fq, o_q) and np.array_equal(res[0], r1.point)
                          and np.array_equal(res[1], r1.dist) and np.array_equal(res[2], r2.point)
                          and np.array_equal(res[3], r2.dist))
    k = len(kept)
    tot = t_bf + t_kd + t_proj
    return {"frames_per_s": k / tot, "ms_per_frame": tot / k * 1e3,
            "ms_per_frame_parts": {"match_features": t_bf / k * 1e3, "kd_build": t_kd / k * 1e3, "projection_pair": t_proj / k * 1e3},
            "sample": f"{k} frames spread over the run (map sizes {[seq.map_size(f) for f in sorted(kept)]})",
            "kind": ("cv2 BFMatcher cross-check (all cores)" if cv2 is not None else "oracle cross-check") +
                    " + oracle/slam_oracle.cpp KD build and projection search (1 core, like the reference)",
            "cores": os.cpu_count(), "gpu_results_identical_on_sample": parity_ok}



def main():
    args = parse_args()
    rank, local_rank, world = dist_env()
    # torchrun exports OMP_NUM_THREADS=1; the CPU arm is meant to use every host core
    os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, local_rank, world)


if __name__ == "__main__":
    main()
