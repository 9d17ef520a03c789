"""Generates tests/golden/triangulate_kitti.npz: inputs and outputs of the OpenCV calls the reference's
Triangulator makes (/root/reference/src/mapping/triangulation.rs:109-114, 161-183): cv2.gemm for P = K [R | t]
and cv2.triangulatePoints on Point2f correspondences.  Run with cv2 4.13.0:  python tests/golden/make_golden_triangulate.py
"""
import os

import cv2
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def scene(n, seed):
    rng = np.random.default_rng(seed)
    K = np.array([[718.856, 0, 607.1928], [0, 718.856, 185.2157], [0, 0, 1.0]])
    R1, t1 = cv2.Rodrigues(np.array([0.002, 0.001, -0.003]))[0], np.array([0.01, -0.02, 0.03])
    R2, t2 = cv2.Rodrigues(np.array([0.01, -0.02, 0.005]))[0], np.array([0.5, 0.02, -0.1])
    X = np.stack([rng.uniform(-12, 12, n), rng.uniform(-3, 3, n), rng.uniform(-8, 60, n)], 1)  # some behind the cameras
    pts = []
    for R, t in ((R1, t1), (R2, t2)):
        c = X @ R.T + t
        pts.append((c[:, :2] / c[:, 2:] * [K[0, 0], K[1, 1]] + [K[0, 2], K[1, 2]] + rng.normal(0, 0.4, (n, 2))).astype(np.float32))
    return K, (R1, t1), (R2, t2), pts[0], pts[1]


def main():
    K, (R1, t1), (R2, t2), p1, p2 = scene(256, 42)
    P = [cv2.gemm(K, np.hstack([R, t[:, None]]), 1.0, None, 0.0) for R, t in ((R1, t1), (R2, t2))]
    h = cv2.triangulatePoints(P[0], P[1], p1.reshape(-1, 1, 2), p2.reshape(-1, 1, 2))
    assert h.dtype == np.float32 and h.shape == (4, 256)
    np.savez_compressed(os.path.join(HERE, "triangulate_kitti.npz"), K=K, R1=R1, t1=t1, R2=R2, t2=t2, p1=p1, p2=p2,
                        P1=P[0], P2=P[1], points4d=h.T.copy(), cv2_version=cv2.__version__)
    print("wrote triangulate_kitti.npz", cv2.__version__)


if __name__ == "__main__":
    main()
