"""Golden vectors for the PnP verification support: cv2 4.13.0's own projectPoints on the
object points of synthetic verify_pnp problems (src/LoopDetector.cpp:188-214), several pose hypotheses each.

    python oracle/gen_golden_pnp.py        # needs cv2 (this container); writes tests/golden/pnp_*.npz

What is pinned: the projection cv::solvePnPRansac's error callback consumes (uv, float32, bit-identical).
The squared error / inlier rule on top of it is recorded twice: from the cv2 projections in numpy float32
(err_np) and asserted equal to the oracle's before anything is written.  TEST INFRASTRUCTURE ONLY."""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import cpu_oracle  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
K4 = np.array([718.856, 718.856, 607.19, 185.22], np.float64)   # fx, fy, cx, cy (KITTI)

CASES = [
    # name, correspondences, hypotheses, outlier fraction, pixel noise, special
    ("small_12", 12, 6, 0.2, 0.5, ""),
    ("typical_300", 300, 40, 0.4, 1.0, ""),
    ("large_4000", 4000, 16, 0.5, 1.5, ""),
    ("threshold_ties", 257, 8, 0.0, 0.0, "ties"),       # errors exactly on / next to the threshold
    ("behind_and_z0", 64, 5, 0.3, 1.0, "degenerate"),   # points behind the camera and on the z = 0 plane
]


def main():
    cpu_oracle.build()
    Kmat = np.array([[K4[0], 0, K4[2]], [0, K4[1], K4[3]], [0, 0, 1]], np.float64)
    for name, n, h, out_frac, noise, special in CASES:
        rng = np.random.default_rng(sum(map(ord, name)))
        obj = (rng.uniform(-8, 8, (n, 3)) + [0, 0, 25]).astype(np.float32)
        rvec0 = rng.uniform(-0.2, 0.2, 3)
        tvec0 = rng.uniform(-0.5, 0.5, 3)
        if special == "degenerate":
            rvec0[:] = 0
            tvec0[:] = 0
            obj[:8, 2] = 0.0                      # z == 0 exactly with R = I, t = 0 (hypothesis 0)
            obj[8:20, 2] = -obj[8:20, 2]          # behind the camera: OpenCV still projects them
        uv0, _ = cv2.projectPoints(obj.reshape(-1, 1, 3), rvec0, tvec0, Kmat, None)
        img = uv0.reshape(-1, 2).astype(np.float32)
        img = np.nan_to_num(img, nan=0.0, posinf=0.0, neginf=0.0)
        img += rng.normal(0, noise, img.shape).astype(np.float32) if noise > 0 else 0
        bad = rng.random(n) < out_frac
        img[bad] = rng.uniform(0, 1241, (int(bad.sum()), 2)).astype(np.float32)
        if special == "ties":
            # image point = projection + (dx, dy) with dx*dx + dy*dy at, just under and just over 16
            off = np.array([[4, 0], [0, 4], [0, -4], [np.nextafter(np.float32(4), np.float32(5)), 0],
                            [np.nextafter(np.float32(4), np.float32(3)), 0], [2.4, 3.2], [3.2, 2.4], [2.828427, 2.828427]],
                           np.float32)
            img = img + off[np.arange(n) % len(off)]
        rvecs = np.stack([rvec0] + [rvec0 + rng.normal(0, 0.0006 * (i % 5), 3) for i in range(1, h)])
        tvecs = np.stack([tvec0] + [tvec0 + rng.normal(0, 0.006 * (i % 5), 3) for i in range(1, h)])
        rt = np.zeros((h, 12), np.float64)
        uv = np.zeros((h, n, 2), np.float32)
        err_np = np.zeros((h, n), np.float32)
        for i in range(h):
            R, _ = cv2.Rodrigues(rvecs[i])
            rt[i, :9] = R.reshape(9)
            rt[i, 9:] = tvecs[i]
            p, _ = cv2.projectPoints(obj.reshape(-1, 1, 3), rvecs[i], tvecs[i], Kmat, None)
            uv[i] = p.reshape(-1, 2)
            with np.errstate(all="ignore"):
                dx = img[:, 0] - uv[i, :, 0]
                dy = img[:, 1] - uv[i, :, 1]
                err_np[i] = (np.float32(0) + dx * dx) + dy * dy
            # the restatement must agree before anything is written
            ouv = cpu_oracle.pnp_project(rt[i], K4, obj)
            assert np.array_equal(ouv.view(np.uint32), uv[i].view(np.uint32)), (name, i)
            c, oerr, omask = cpu_oracle.pnp_score(rt[i], K4, obj, img, 4.0)
            assert np.array_equal(oerr.view(np.uint32), err_np[i].view(np.uint32)), (name, i)
            with np.errstate(all="ignore"):
                assert np.array_equal(omask, (err_np[i] <= np.float32(16.0)).astype(np.uint8)) and c == int(omask.sum())
        np.savez_compressed(os.path.join(GOLD, f"pnp_{name}.npz"), object=obj, image=img, rvec=rvecs, tvec=tvecs,
                            rt=rt, K4=K4, uv=uv, err_np=err_np)
        print(name, "inliers per hypothesis:", (err_np <= 16).sum(1)[:8])


if __name__ == "__main__":
    main()
