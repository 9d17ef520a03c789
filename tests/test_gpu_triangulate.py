"""GPU: csrc/triangulate.cu through the C ABI vs the oracle (oracle/restated.py::triangulate, pinned
to cv2.triangulatePoints in tests/test_triangulation_cpu.py), plus ports of the reference's own Triangulator tests
(/root/reference/src/mapping/triangulation.rs:200-355).

Tolerance (floating point): the homogeneous point comes out of an SVD.  The reference's is LAPACK's through OpenCV,
ours is a one-sided Jacobi in f64; both are rounded to fp32 (the CV_32F the reference reads), so the bar is
>= 99.9 % of well-conditioned points bit-identical in fp32 (up to the SVD's arbitrary sign) and every one within
4 fp32 ulp relative to the vector's norm (1); kept sets and positions must then be identical wherever the fp32
vectors are.
"""
import os

import numpy as np
import pytest

from oracle import restated as R

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "triangulate_kitti.npz")
KITTI = (718.856, 718.856, 607.1928, 185.2157)


@pytest.fixture(scope="module")
def orbx():
    import slam_rs_b200 as s
    return s


def sign_aligned(a, b):
    s = np.where((a * b).sum(1) < 0, -1.0, 1.0).astype(np.float32)[:, None]
    return a * s


def scene(n, seed, noise=0.4):
    rng = np.random.default_rng(seed)
    from tests.test_gpu_map import rot
    R1, t1 = rot(0.002, 0.001, -0.003), np.array([0.01, -0.02, 0.03])
    R2, t2 = rot(0.01, -0.02, 0.005), np.array([0.5, 0.02, -0.1])
    X = np.stack([rng.uniform(-12, 12, n), rng.uniform(-3, 3, n), rng.uniform(-8, 60, n)], 1)
    pts = []
    for Rm, t in ((R1, t1), (R2, t2)):
        c = X @ Rm.T + t
        pts.append((c[:, :2] / c[:, 2:] * [KITTI[0], KITTI[1]] + [KITTI[2], KITTI[3]]
                    + rng.normal(0, noise, (n, 2))).astype(np.float32))
    return (R1, t1), (R2, t2), pts[0], pts[1]


def check_against_oracle(orbx, tri, pose1, pose2, p1, p2, desc=None):
    ids, xyz, od, hd, p4 = tri.triangulate(pose1, pose2, p1, p2, desc, return_points4d=True)
    P1 = R.projection_matrix(KITTI, *pose1)
    P2 = R.projection_matrix(KITTI, *pose2)
    want4 = R.triangulate_points4d(P1, P2, p1, p2)
    got4 = sign_aligned(p4, want4)
    same = np.all(got4 == want4, axis=1)
    assert same.mean() >= 0.999, same.mean()
    assert np.abs(got4.astype(np.float64) - want4).max() <= 4 * 2.0 ** -24
    wi, wx, wd, wh = R.triangulate(KITTI, pose1, pose2, p1, p2, desc)
    # kept set: identical on the points whose fp32 vector is identical
    keep_got = np.zeros(len(p1), bool)
    keep_got[ids] = True
    keep_want = np.zeros(len(p1), bool)
    keep_want[wi] = True
    assert np.array_equal(keep_got[same], keep_want[same])
    assert np.all(np.diff(ids) > 0)
    both = np.intersect1d(ids[same[ids]], wi)
    gx = xyz[np.searchsorted(ids, both)]
    assert np.array_equal(gx, wx[np.searchsorted(wi, both)])                   # f64 positions bit-identical
    rows = 0 if desc is None else len(desc)
    assert np.array_equal(hd, ids < rows)
    if rows:
        assert np.array_equal(od[hd], np.asarray(desc)[ids[hd]]) and not od[~hd].any()
    return ids, xyz, same.mean()


def test_golden_vector(orbx):
    g = np.load(GOLD)
    tri = orbx.Triangulator(orbx.CameraIntrinsics.kitti(), max_points=1024)
    ids, xyz, od, hd, p4 = tri.triangulate((g["R1"], g["t1"]), (g["R2"], g["t2"]), g["p1"], g["p2"], None,
                                           return_points4d=True)
    want = g["points4d"]                                                        # cv2.triangulatePoints itself
    got = sign_aligned(p4, want)
    assert np.all(got == want, axis=1).mean() >= 0.999
    assert np.abs(got.astype(np.float64) - want).max() <= 4 * 2.0 ** -24
    assert 0 < len(ids) < 256 and not hd.any() and not od.any()


@pytest.mark.parametrize("n,seed", [(1, 0), (31, 1), (1025, 2), (5000, 3)])
def test_vs_oracle(orbx, n, seed):
    pose1, pose2, p1, p2 = scene(n, seed)
    tri = orbx.Triangulator(orbx.CameraIntrinsics.kitti(), max_points=8192)
    rng = np.random.default_rng(seed)
    desc = rng.integers(0, 256, (max(n // 2, 1), 32), dtype=np.uint8)          # fewer rows than points (:142)
    check_against_oracle(orbx, tri, pose1, pose2, p1, p2, desc)
    check_against_oracle(orbx, tri, pose1, pose2, p1, p2, None)


def test_reference_triangulator_tests(orbx):
    cam = orbx.CameraIntrinsics.kitti()
    tri = orbx.Triangulator(cam, max_points=64)
    assert tri.min_parallax_deg == 1.0 and tri.max_reproj_error == 4.0         # test_triangulator_creation (:243-248)
    tri2 = orbx.Triangulator(cam, max_points=64).with_min_parallax(2.0).with_max_reproj_error(5.0)
    assert tri2.min_parallax_deg == 2.0 and tri2.max_reproj_error == 5.0       # test_triangulator_builder (:251-258)
    pose1 = (np.eye(3), np.zeros(3))
    pose2 = (np.eye(3), np.array([1.0, 0.0, 0.0]))
    ids, xyz, od, hd = tri.triangulate(pose1, pose2, np.zeros((0, 2)), np.zeros((0, 2)))
    assert len(ids) == 0 and xyz.shape == (0, 3)                               # test_triangulate_empty_points (:261-275)
    with pytest.raises(orbx.OrbxError, match="same length"):                   # test_triangulate_mismatched_points (:278-293)
        tri.triangulate(pose1, pose2, [[100.0, 100.0]], np.zeros((0, 2)))
    cam2 = orbx.CameraIntrinsics(500.0, 500.0, 320.0, 240.0)                   # test_triangulate_synthetic_points (:311-354)
    tri3 = orbx.Triangulator(cam2, max_points=64)
    X = np.array([[0.0, 0.0, 10.0], [2.0, 1.0, 10.0]])
    pts = []
    for Rm, t in (pose1, pose2):
        c = X @ Rm.T + t
        pts.append((c[:, :2] / c[:, 2:] * 500.0 + [320.0, 240.0]).astype(np.float32))
    ids, xyz, _, _ = tri3.triangulate(pose1, pose2, pts[0], pts[1])
    assert ids.tolist() == [0, 1]
    assert np.all((xyz[:, 2] > 5.0) & (xyz[:, 2] < 15.0)) and np.allclose(xyz, X, atol=1e-3)
    behind = tri3.triangulate(pose1, pose2, pts[1], pts[0])                    # swapped views: negative depth
    assert len(behind[0]) == 0                                                 # test_is_in_front_of_camera (:296-308)
    with pytest.raises(orbx.OrbxError):
        tri.triangulate(pose1, pose2, np.zeros((65, 2)), np.zeros((65, 2)))    # > max_points


def test_to_world_and_map_feed(orbx):
    """point_cloud.rs:261-279: triangulate -> transform_points_to_world -> Map::add_points -> find_matches."""
    pose1, pose2, p1, p2 = scene(3000, 7)
    cam = orbx.CameraIntrinsics.kitti()
    tri = orbx.Triangulator(cam, max_points=4096)
    rng = np.random.default_rng(7)
    desc = rng.integers(0, 256, (3000, 32), dtype=np.uint8)
    ids, xyz, od, hd = tri.triangulate(pose1, pose2, p1, p2, desc)
    world = tri.transform_points_to_world(xyz, pose2)
    assert np.array_equal(world, R.transform_points_to_world(xyz, pose2))      # unfused f64, nalgebra's order
    m = orbx.Map(cam, capacity=4096)
    got_ids = m.add_points(world, od, has_descriptor=hd)
    assert m.size() == len(ids) and got_ids.tolist() == list(range(len(ids)))
    matcher = orbx.FeatureMatcher()
    pairs = m.find_matches(None, desc, pose2, matcher)
    want = R.map_find_matches(world, od, hd, np.ones(len(ids), bool), pose2[0], pose2[1], KITTI, desc)
    assert np.array_equal(pairs, want) and len(pairs) > 0


def test_device_pointers(orbx):
    import ctypes as C
    import torch
    from slam_rs_b200._ffi import lib, check
    pose1, pose2, p1, p2 = scene(2000, 11)
    tri = orbx.Triangulator(orbx.CameraIntrinsics.kitti(), max_points=2048)
    h_ids, h_xyz, _, _ = tri.triangulate(pose1, pose2, p1, p2)
    dev = torch.device("cuda", 0)
    d1, d2 = torch.from_numpy(p1).to(dev), torch.from_numpy(p2).to(dev)
    ids = torch.zeros(2000, dtype=torch.int32, device=dev)
    xyz = torch.zeros((2000, 3), dtype=torch.float64, device=dev)
    n = C.c_int(0)
    arrs = [np.ascontiguousarray(a, np.float64) for a in (pose1[0], pose1[1], pose2[0], pose2[1])]
    ptr = lambda a: C.c_void_p(a.ctypes.data)
    check(lib.orbx_triangulate(tri._h, ptr(arrs[0]), ptr(arrs[1]), ptr(arrs[2]), ptr(arrs[3]),
                               C.c_void_p(d1.data_ptr()), 2000, C.c_void_p(d2.data_ptr()), 2000, None, 0,
                               C.c_void_p(ids.data_ptr()), C.c_void_p(xyz.data_ptr()), None, None, None, 2000,
                               C.byref(n)))
    assert n.value == len(h_ids)
    assert np.array_equal(ids[:n.value].cpu().numpy(), h_ids) and np.array_equal(xyz[:n.value].cpu().numpy(), h_xyz)


def test_large_batch_properties(orbx):
    """200 k correspondences: every kept point is in front of both cameras and reprojects onto its observations."""
    n = 200_000
    pose1, pose2, p1, p2 = scene(n, 21, noise=0.0)
    tri = orbx.Triangulator(orbx.CameraIntrinsics.kitti(), max_points=n)
    ids, xyz, _, _ = tri.triangulate(pose1, pose2, p1, p2)
    assert len(ids) > n // 2 and np.all(np.diff(ids) > 0)
    for (Rm, t), p in ((pose1, p1), (pose2, p2)):
        c = xyz @ Rm.T + t
        assert np.all(c[:, 2] > 0)
        uv = c[:, :2] / c[:, 2:] * [KITTI[0], KITTI[1]] + [KITTI[2], KITTI[3]]
        err = np.linalg.norm(uv - p[ids], axis=1)
        assert np.median(err) < 1e-2


def test_sequence_front_end_triangulates_every_pair(orbx):
    """SequenceFrontEnd(pose=..., triangulate=True): the batched device form (orbx_triangulate_batch, poses read from the
    pose stage's device output) returns, for every pair, exactly what the single call returns for the same inputs."""
    import torch
    from oracle import synth
    from slam_rs_b200.frontend import SequenceFrontEnd
    T, h, w, nf = 7, 376, 1241, 800
    seq = synth.sequence(h, w, T, canvas_seed=41, shift=24)
    cam = orbx.CameraIntrinsics.kitti()
    single = orbx.Triangulator(cam, max_points=4096)
    for kwargs in ({"chunk": T}, {"chunk": T, "host_chunk": 3, "dual": True}):
        fe = SequenceFrontEnd(w, h, nfeatures=nf, max_frames=T, pose=cam, triangulate=True, **kwargs)
        out = None
        if "dual" in kwargs:
            out = fe.alloc_host_outputs()
            fe.run_host(torch.from_numpy(seq).pin_memory(), out)
        else:
            dev = torch.zeros((T, h, fe.pitch), dtype=torch.uint8, device="cuda")
            dev[:, :, :w] = torch.from_numpy(seq).cuda()
            fe.run_device(dev)
        torch.cuda.synchronize()
        total = 0
        for t in range(1, T):
            p1, p2 = fe.pair_points(t)
            _, Rm, tv, _, _, _ = fe.pair_pose(t)
            ids, xyz = fe.pair_map_points(t)
            wi, wx, _, _ = single.triangulate((np.eye(3), np.zeros(3)), (Rm, tv), p1, p2)
            assert np.array_equal(ids, wi) and np.array_equal(xyz, wx)
            if out is not None:                                                # the host copies run_host made
                m = int(out["tri_n"][t])
                assert m == len(ids) and np.array_equal(out["tri_ids"][t, :m].numpy(), ids)
                assert np.array_equal(out["tri_xyz"][t, :m].numpy(), xyz)
            total += len(ids)
        assert total > 0
