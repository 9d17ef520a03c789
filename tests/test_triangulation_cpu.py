"""CPU: the oracle's restatement of mapping::Triangulator (oracle/restated.py::triangulate) pinned
against the OpenCV calls the reference makes (/root/reference/src/mapping/triangulation.rs:109-114, 161-183) — the
committed golden vector and cv2 live — plus the closed-form checks of the reference's own tests (:200-355)."""
import os

import numpy as np
import pytest

from oracle import restated as R

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "triangulate_kitti.npz")
KITTI = (718.856, 718.856, 607.1928, 185.2157)


def same_up_to_sign(a, b):
    s = np.where((a * b).sum(1) < 0, -1.0, 1.0).astype(np.float32)[:, None]
    return np.all(a * s == b, axis=1)


def test_projection_and_points4d_vs_golden():
    g = np.load(GOLD)
    P1 = R.projection_matrix(KITTI, g["R1"], g["t1"])
    P2 = R.projection_matrix(KITTI, g["R2"], g["t2"])
    assert np.array_equal(P1, g["P1"]) and np.array_equal(P2, g["P2"])          # == cv::gemm, bit for bit
    h = R.triangulate_points4d(P1, P2, g["p1"], g["p2"])
    assert h.dtype == np.float32 and same_up_to_sign(h, g["points4d"]).all()   # == cv::triangulatePoints (CV_32F)


def test_points4d_vs_cv2_live():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(3)
    for trial in range(3):
        n = 500
        R2 = cv2.Rodrigues(rng.normal(0, 0.03, 3))[0]
        t2 = rng.normal(0, 0.4, 3)
        X = np.stack([rng.uniform(-10, 10, n), rng.uniform(-3, 3, n), rng.uniform(-5, 50, n)], 1)
        K = np.array([[KITTI[0], 0, KITTI[2]], [0, KITTI[1], KITTI[3]], [0, 0, 1.0]])
        pts = []
        for Rm, t in ((np.eye(3), np.zeros(3)), (R2, t2)):
            c = X @ Rm.T + t
            pts.append((c[:, :2] / c[:, 2:] * [K[0, 0], K[1, 1]] + [K[0, 2], K[1, 2]]).astype(np.float32))
        P1 = cv2.gemm(K, np.hstack([np.eye(3), np.zeros((3, 1))]), 1.0, None, 0.0)
        P2 = cv2.gemm(K, np.hstack([R2, t2[:, None]]), 1.0, None, 0.0)
        assert np.array_equal(P1, R.projection_matrix(KITTI, np.eye(3), np.zeros(3)))
        assert np.array_equal(P2, R.projection_matrix(KITTI, R2, t2))
        want = cv2.triangulatePoints(P1, P2, pts[0].reshape(-1, 1, 2), pts[1].reshape(-1, 1, 2)).T
        assert same_up_to_sign(R.triangulate_points4d(P1, P2, pts[0], pts[1]), want).all()


def test_reference_triangulator_tests():
    """triangulation.rs:258-355: empty -> empty, mismatched -> error, two synthetic points at depth 10."""
    pose1 = (np.eye(3), np.zeros(3))
    pose2 = (np.eye(3), np.array([1.0, 0.0, 0.0]))
    ids, xyz, desc, has = R.triangulate(KITTI, pose1, pose2, np.zeros((0, 2)), np.zeros((0, 2)))
    assert len(ids) == 0 and xyz.shape == (0, 3)
    with pytest.raises(ValueError):
        R.triangulate(KITTI, pose1, pose2, [[100.0, 100.0]], np.zeros((0, 2)))
    intr = (500.0, 500.0, 320.0, 240.0)
    X = np.array([[0.0, 0.0, 10.0], [2.0, 1.0, 10.0]])
    pts = []
    for Rm, t in (pose1, pose2):
        c = X @ Rm.T + t
        pts.append((c[:, :2] / c[:, 2:] * 500.0 + [320.0, 240.0]).astype(np.float32))
    ids, xyz, _, has = R.triangulate(intr, pose1, pose2, pts[0], pts[1])
    assert ids.tolist() == [0, 1] and not has.any()
    assert np.all((xyz[:, 2] > 5.0) & (xyz[:, 2] < 15.0)) and np.allclose(xyz, X, atol=1e-3)


def test_cheirality_and_descriptor_attach():
    g = np.load(GOLD)
    d = np.arange(100 * 32, dtype=np.uint8).reshape(100, 32)                   # fewer rows than points (:142)
    ids, xyz, desc, has = R.triangulate(KITTI, (g["R1"], g["t1"]), (g["R2"], g["t2"]), g["p1"], g["p2"], d)
    assert 0 < len(ids) < 256 and np.all(np.diff(ids) > 0)                     # some points are behind a camera
    assert np.array_equal(has, ids < 100) and np.array_equal(desc[has], d[ids[has]]) and not desc[~has].any()
    for Rm, t in ((g["R1"], g["t1"]), (g["R2"], g["t2"])):
        assert np.all((xyz @ Rm.T + t)[:, 2] > 0)


def test_transform_points_to_world_roundtrip():
    g = np.load(GOLD)
    rng = np.random.default_rng(0)
    p = rng.normal(0, 5, (50, 3))
    w = R.transform_points_to_world(p, (g["R2"], g["t2"]))
    assert np.allclose(w @ g["R2"].T + g["t2"], p, atol=1e-12)
