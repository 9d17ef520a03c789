"""Generates tests/golden/pose_kitti.npz: inputs and outputs of the OpenCV calls the reference's PoseEstimator makes
(/root/reference/src/odometry/pose.rs:64-73 find_essential_mat(..., RANSAC, 0.999, 1.0, 1000) and pose.rs:94-104
recover_pose_estimated) on seeded two-view scenes with the KITTI intrinsics.  Run with cv2 4.13.0:
    python tests/golden/make_golden_pose.py
"""
import os
import sys

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from tests.test_pose_cpu import two_view_scene, K  # noqa: E402  (the scene generator the tests use)

CASES = [(300, 21, 0.2), (1000, 22, 0.35), (40, 23, 0.1)]       # (points, seed, outlier share)


def main():
    out = {"K": K, "cv2_version": cv2.__version__, "cases": np.array(CASES, np.float64)}
    for i, (n, seed, outl) in enumerate(CASES):
        p1, p2, R, t = two_view_scene(n, seed, outliers=outl)
        E, m = cv2.findEssentialMat(p1, p2, K, cv2.RANSAC, 0.999, 1.0, 1000)
        assert E.shape == (3, 3)
        good, Rr, tr, pm = cv2.recoverPose(E, p1, p2, K, mask=m.copy())
        good0, R0, t0, pm0 = cv2.recoverPose(E, p1, p2, K)          # the reference passes no input mask (pose.rs:92, 103)
        out.update({f"p1_{i}": p1, f"p2_{i}": p2, f"R_true_{i}": R, f"t_true_{i}": t, f"E_{i}": E,
                    f"mask_{i}": m.ravel().astype(np.uint8), f"R_{i}": R0, f"t_{i}": t0.ravel(),
                    f"pose_mask_{i}": (pm0.ravel() != 0).astype(np.uint8), f"good_{i}": np.int64(good0),
                    f"good_masked_{i}": np.int64(good)})
    np.savez_compressed(os.path.join(HERE, "pose_kitti.npz"), **out)
    print("wrote pose_kitti.npz", cv2.__version__)


if __name__ == "__main__":
    main()
