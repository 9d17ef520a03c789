"""CPU: the oracle's restatement of mapping::Map's hot path (oracle/restated.py::map_project / map_find_matches)
against the closed-form checks the reference's own tests make (/root/reference/src/mapping/map.rs:183-240)."""
import numpy as np

from oracle import restated as R, synth

KITTI = (718.856, 718.856, 607.1928, 185.2157)      # CameraIntrinsics::kitti() (camera.rs:19-26)


def test_projection_principal_point():
    """map.rs:207-221 test_projection: (0,0,10) under identity pose projects to (cx, cy)."""
    xy, vis = R.map_project([[0.0, 0.0, 10.0]], np.eye(3), np.zeros(3), *KITTI)
    assert vis[0] and abs(xy[0, 0] - KITTI[2]) < 1e-6 and abs(xy[0, 1] - KITTI[3]) < 1e-6


def test_projection_cull_rules():
    pts = np.array([[0, 0, -1.0], [0, 0, 0.0], [1e3, 0, 1.0], [-1.0, 0, 1.0], [0, 5.0, 1.0], [0.5, 0.2, 2.0]])
    xy, vis = R.map_project(pts, np.eye(3), np.zeros(3), *KITTI)
    # behind / on the camera plane (map.rs:61), x >= 4000, x < 0, y >= 3000 (map.rs:90), inside
    assert vis.tolist() == [False, False, False, False, False, True]


def test_find_matches_semantics():
    rng = np.random.default_rng(5)
    mp, fr = synth.map_and_frame_descriptors(500, 60, f=3)
    pts = np.stack([rng.uniform(-5, 5, 500), rng.uniform(-2, 2, 500), rng.uniform(-3, 30, 500)], 1)
    has = rng.random(500) < 0.9
    alive = np.ones(500, bool)
    out = R.map_find_matches(pts, mp, has, alive, np.eye(3), np.zeros(3), KITTI, fr)
    _, vis = R.map_project(pts, np.eye(3), np.zeros(3), *KITTI)
    ids = np.nonzero(vis & has)[0]
    assert len(out) and set(out[:, 0]) <= set(ids) and np.all(np.diff(out[:, 0]) > 0)
    # every reported pair is the first-min train row of that map point and passes the min-distance filter
    D = R.hamming_matrix(mp[ids], fr)
    thr = max(2.0 * D.min(axis=1).min(), 30.0)
    want = [(i, int(D[k].argmin())) for k, i in enumerate(ids) if D[k].min() < thr]
    assert [tuple(r) for r in out.tolist()] == want
    assert len(R.map_find_matches(pts, mp, has, alive, np.eye(3), np.zeros(3), KITTI, fr[:0])) == 0


def test_extract_matched_points_restatement():
    """oracle/restated.py::extract_matched_points == pose.rs:27-44 (a plain gather; out of range -> error)."""
    import pytest
    kp1 = np.arange(10, dtype=np.float32).reshape(5, 2)
    kp2 = 100 + np.arange(8, dtype=np.float32).reshape(4, 2)
    m = np.array([[4, 0], [0, 3], [2, 2]])
    p1, p2 = R.extract_matched_points(kp1, kp2, m)
    assert np.array_equal(p1, kp1[[4, 0, 2]]) and np.array_equal(p2, kp2[[0, 3, 2]])
    e1, e2 = R.extract_matched_points(kp1, kp2, m[:0])
    assert e1.shape == (0, 2) and e2.shape == (0, 2)
    with pytest.raises(IndexError):
        R.extract_matched_points(kp1, kp2, np.array([[5, 0]]))
    with pytest.raises(IndexError):
        R.extract_matched_points(kp1, kp2, np.array([[0, -1]]))


def test_camera_intrinsics_mirror():
    """camera.rs:58-85 (test_camera_creation, test_kitti_preset, test_to_matrix) over the Python mirror."""
    from slam_rs_b200.mapping import CameraIntrinsics
    cam = CameraIntrinsics(500.0, 500.0, 320.0, 240.0)
    assert (cam.fx, cam.fy, cam.cx, cam.cy) == (500.0, 500.0, 320.0, 240.0)
    k = CameraIntrinsics.kitti()
    assert k.fx > 700.0 and k.fy > 700.0
    v = CameraIntrinsics.webcam_vga()
    assert (v.fx, v.cx, v.cy) == (500.0, 320.0, 240.0)
    m = CameraIntrinsics(100.0, 200.0, 50.0, 75.0).to_matrix()
    assert m.shape == (3, 3) and m.dtype == np.float64
    assert np.array_equal(m, [[100.0, 0, 50.0], [0, 200.0, 75.0], [0, 0, 1.0]])
