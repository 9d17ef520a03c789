"""GPU: csrc/pose.cu (PoseEstimator::compute_essential_matrix / recover_pose, pose.rs:47-114) through
the C ABI vs cv2.findEssentialMat / cv2.recoverPose — the OpenCV functions the reference calls (cv2 is the oracle here;
its numpy restatement is pinned in tests/test_pose_cpu.py).

Floating point, so the bar is a tolerance: the kernels replay OpenCV's RANSAC draws, so on every scene the SAME
hypothesis must win — E equal up to sign to 1e-6 (different five-point arithmetic: Gauss-Jordan + bisection here, SVD +
companion eigenvalues there), identical inlier masks except points whose fp32 Sampson error sits within 1e-4 relative
of the threshold, R / t within 1e-6.
"""
import numpy as np
import pytest

from oracle import pose_restated as PR
from tests.test_pose_cpu import two_view_scene, golden_pose_cases, K, KITTI

pytestmark = pytest.mark.gpu
cv2 = pytest.importorskip("cv2")


@pytest.fixture(scope="module")
def orbx():
    import slam_rs_b200 as s
    return s


def e_diff(a, b):
    return min(np.abs(a - b).max(), np.abs(a + b).max())


@pytest.mark.parametrize("n,seed,outliers", [(500, 1, 0.25), (2000, 2, 0.3), (64, 3, 0.1), (1500, 4, 0.5), (9, 5, 0.0)])
def test_essential_matrix_vs_cv2(orbx, n, seed, outliers):
    p1, p2, _, _ = two_view_scene(n, seed, outliers=outliers)
    pe = orbx.PoseEstimator(orbx.CameraIntrinsics.kitti(), max_points=4096)
    E = pe.compute_essential_matrix(p1, p2)
    Ecv, mcv = cv2.findEssentialMat(p1, p2, K, cv2.RANSAC, 0.999, 1.0, 1000)
    assert Ecv.shape == (3, 3)
    assert e_diff(E, Ecv) < 1e-6, (E, Ecv)
    mcv = mcv.ravel() != 0
    assert abs(np.linalg.norm(E) - 1.0) < 1e-12
    assert (pe.last_mask != mcv).sum() <= max(1, n // 500)
    assert abs(pe.last_inliers - int(mcv.sum())) <= max(1, n // 500)
    assert 0 < pe.last_iterations <= 1000


@pytest.mark.parametrize("seed", [6, 7, 8])
def test_recover_pose_vs_cv2(orbx, seed):
    p1, p2, Rt, tt = two_view_scene(800, seed, outliers=0.15)
    Ecv, _ = cv2.findEssentialMat(p1, p2, K, cv2.RANSAC, 0.999, 1.0, 1000)
    good, Rcv, tcv, mcv = cv2.recoverPose(Ecv, p1, p2, K)
    pe = orbx.PoseEstimator(orbx.CameraIntrinsics.kitti(), max_points=1024)
    R, t = pe.recover_pose(Ecv, p1, p2)
    assert np.abs(R - Rcv).max() < 1e-6 and np.abs(t - tcv.ravel()).max() < 1e-6
    assert pe.last_good == good and np.array_equal(pe.last_pose_mask, mcv.ravel() != 0)
    assert np.abs(R - Rt).max() < 0.02 and abs(np.linalg.det(R) - 1.0) < 1e-9


def test_committed_golden(orbx):
    """tests/golden/pose_kitti.npz: OpenCV's outputs committed as a fixture (tests/golden/make_golden_pose.py)."""
    pe = orbx.PoseEstimator(orbx.CameraIntrinsics.kitti(), max_points=1024)
    for c in golden_pose_cases():
        n = len(c["p1"])
        E = pe.compute_essential_matrix(c["p1"], c["p2"])
        assert e_diff(E, c["E"]) < 1e-6
        assert (pe.last_mask != (c["mask"] != 0)).sum() <= max(1, n // 500)
        R, t = pe.recover_pose(c["E"], c["p1"], c["p2"])
        assert np.abs(R - c["R"]).max() < 1e-6 and np.abs(t - c["t"]).max() < 1e-6
        assert pe.last_good == int(c["good"]) and np.array_equal(pe.last_pose_mask, c["pose_mask"] != 0)


def test_full_chain_like_visual_odometry(orbx):
    """visual_odometry.rs:111-117: matched points -> E -> (R, t), from our own E."""
    p1, p2, Rt, tt = two_view_scene(1200, 9, outliers=0.2)
    pe = orbx.PoseEstimator(orbx.CameraIntrinsics.kitti(), max_points=2048)
    E = pe.compute_essential_matrix(p1, p2)
    R, t = pe.recover_pose(E, p1, p2)
    Ecv, _ = cv2.findEssentialMat(p1, p2, K, cv2.RANSAC, 0.999, 1.0, 1000)
    _, Rcv, tcv, _ = cv2.recoverPose(Ecv, p1, p2, K)
    assert np.abs(R - Rcv).max() < 1e-5 and np.abs(t - tcv.ravel()).max() < 1e-5
    tn = tt / np.linalg.norm(tt)
    assert np.abs(R - Rt).max() < 0.02 and np.abs(t - tn).max() < 0.05


def test_reference_error_cases(orbx):
    pe = orbx.PoseEstimator(orbx.CameraIntrinsics.kitti(), max_points=256)
    p = np.zeros((7, 2), np.float32)
    with pytest.raises(orbx.OrbxError, match="Insufficient points: 7"):        # pose.rs:52-59
        pe.compute_essential_matrix(p, p)
    # degenerate input: all points identical -> cv2 returns a model or none; ours must not crash and must say which
    q = np.tile(np.array([[100.0, 100.0]], np.float32), (20, 1))
    try:
        pe.compute_essential_matrix(q, q)
    except orbx.OrbxError as e:
        assert e.code == -5
    # recover_pose with points that cannot be in front of any candidate: "Too few inliers" (pose.rs:104-106)
    p1, p2, _, _ = two_view_scene(200, 10, outliers=0.0)
    Ecv, _ = cv2.findEssentialMat(p1, p2, K, cv2.RANSAC, 0.999, 1.0, 1000)
    with pytest.raises(orbx.OrbxError, match="Too few inliers"):
        pe.recover_pose(Ecv, p1[:4], p2[:4])
    with pytest.raises(orbx.OrbxError):
        pe.compute_essential_matrix(np.zeros((300, 2), np.float32), np.zeros((300, 2), np.float32))   # > max_points


def test_batch_device_matches_single_calls(orbx):
    import torch
    npairs, rows = 12, 700
    cam = orbx.CameraIntrinsics.kitti()
    pe = orbx.PoseEstimator(cam, max_points=rows, max_problems=npairs)
    single = orbx.PoseEstimator(cam, max_points=rows)
    P1 = np.zeros((npairs, rows, 2), np.float32)
    P2 = np.zeros((npairs, rows, 2), np.float32)
    ns = np.zeros(npairs, np.int32)
    for p in range(npairs):
        n = [650, 300, 5, 700, 40][p % 5]                                      # one problem has too few points
        a, b, _, _ = two_view_scene(n, 100 + p, outliers=0.2)
        P1[p, :n], P2[p, :n], ns[p] = a, b, n
    dev = torch.device("cuda", 0)
    d1, d2, dn = torch.from_numpy(P1).to(dev), torch.from_numpy(P2).to(dev), torch.from_numpy(ns).to(dev)
    E = torch.zeros((npairs, 3, 3), dtype=torch.float64, device=dev)
    R = torch.zeros((npairs, 3, 3), dtype=torch.float64, device=dev)
    t = torch.zeros((npairs, 3), dtype=torch.float64, device=dev)
    ni = torch.zeros(npairs, dtype=torch.int32, device=dev)
    ng = torch.zeros(npairs, dtype=torch.int32, device=dev)
    st = torch.zeros(npairs, dtype=torch.int32, device=dev)
    rm = torch.zeros((npairs, rows), dtype=torch.uint8, device=dev)
    pe.estimate_pose_batch_device(d1, d2, dn, npairs, rows, E, R, t, ni, ng, st, rm, None)
    torch.cuda.synchronize()
    for p in range(npairs):
        n = int(ns[p])
        if n < 8:
            assert int(st[p]) == 1
            continue
        Es = single.compute_essential_matrix(P1[p, :n], P2[p, :n])
        Rs, ts = single.recover_pose(Es, P1[p, :n], P2[p, :n])
        assert int(st[p]) == 0
        assert np.array_equal(E[p].cpu().numpy(), Es) and int(ni[p]) == single.last_inliers
        assert np.array_equal(rm[p, :n].cpu().numpy() != 0, single.last_mask)
        assert np.array_equal(R[p].cpu().numpy(), Rs) and np.array_equal(t[p].cpu().numpy(), ts)
        assert int(ng[p]) == single.last_good


def test_sequence_front_end_pose(orbx):
    """SequenceFrontEnd(pose=intrinsics): frames -> keypoints -> matches -> matched points -> E -> (R, t) without leaving
    the device; every pair agrees with cv2 run on the same matched points (both pipeline forms)."""
    import torch
    from oracle import synth
    from slam_rs_b200.frontend import SequenceFrontEnd
    T, h, w, nf = 9, 376, 1241, 1000
    # 24 px of image shift per frame: depth f / 24 = 30 baselines, inside cv::recoverPose's distance threshold of 50
    # (the 3 px / frame of the benchmark sequence puts every point at 240 and OpenCV, like us, keeps none)
    seq = synth.sequence(h, w, T, canvas_seed=31, shift=24)
    cam = orbx.CameraIntrinsics.kitti()
    same = pairs = 0
    # pose_chunk=3: the pairs of a chunk go through the pose stage in several launch sets (the default takes them in one)
    for kwargs in ({"chunk": T, "pose_chunk": 3}, {"chunk": T, "host_chunk": 3, "dual": True}):
        fe = SequenceFrontEnd(w, h, nfeatures=nf, max_frames=T, pose=cam, **kwargs)
        if "dual" in kwargs:
            out = fe.alloc_host_outputs()
            fe.run_host(torch.from_numpy(seq).pin_memory(), out)
        else:
            dev = torch.zeros((T, h, fe.pitch), dtype=torch.uint8, device="cuda")
            dev[:, :, :w] = torch.from_numpy(seq).cuda()
            fe.run_device(dev)
        torch.cuda.synchronize()
        if "dual" in kwargs:                                                   # run_host also brings the poses back
            for key in ("pose_E", "pose_R", "pose_t", "pose_inliers", "pose_good", "pose_status"):
                assert np.array_equal(out[key][1:].numpy(), getattr(fe, key)[1:].cpu().numpy()), key
            assert fe.d2h_bytes(T) == sum(v.numel() * v.element_size() for v in out.values())
        for t in range(1, T):
            p1, p2 = fe.pair_points(t)
            E, R, tv, ninl, good, status = fe.pair_pose(t)
            assert status == 0 and len(p1) >= 8, (status, len(p1), good)
            Ecv, mcv = cv2.findEssentialMat(p1, p2, K, cv2.RANSAC, 0.999, 1.0, 1000)
            # The synthetic scene is ONE fronto-parallel plane: the degenerate case of the five-point problem.  Every
            # sample has two valid solutions (the planar two-fold ambiguity) that explain all points equally well, the
            # first one in solver order keeps the lead (strict >), and OpenCV's solution order (that of cv::solvePoly's
            # roots) is not reproduced.  So: either the same model (E within 1e-4, conditioning shows) and then the same
            # pose, or a tie — our model has as many inliers under the oracle's own error function as cv2's has.
            ncv = int(mcv.sum())
            assert abs(ninl - ncv) <= max(2, len(p1) // 100)
            if e_diff(E, Ecv) < 1e-4:
                same += 1
                g, Rcv, tcv, _ = cv2.recoverPose(Ecv, p1, p2, K)
                assert np.abs(R - Rcv).max() < 1e-3 and np.abs(tv - tcv.ravel()).max() < 1e-3
                assert abs(good - g) <= max(2, len(p1) // 100)
            else:
                x1, x2 = PR.normalize_points(p1, KITTI), PR.normalize_points(p2, KITTI)
                thr = np.float32((1.0 / KITTI[0]) ** 2)
                assert int((PR.sampson_errors(E, x1, x2) <= thr).sum()) >= ncv - max(2, len(p1) // 100)
                g, Rm, tm, _ = cv2.recoverPose(E, p1, p2, K)                    # our E through OpenCV's recoverPose
                assert np.abs(R - Rm).max() < 1e-3 and np.abs(tv - tm.ravel()).max() < 1e-3
            pairs += 1
    assert same >= pairs // 2, (same, pairs)
    # the benchmark sequence itself (3 px / frame): every point is farther than 50 baselines, so cv::recoverPose keeps
    # none and the reference reports "Too few inliers" (pose.rs:104-106) — same here
    seq3 = synth.sequence(h, w, 3, canvas_seed=31)
    fe = SequenceFrontEnd(w, h, nfeatures=nf, max_frames=3, pose=cam, chunk=3)
    dev = torch.zeros((3, h, fe.pitch), dtype=torch.uint8, device="cuda")
    dev[:, :, :w] = torch.from_numpy(seq3).cuda()
    fe.run_device(dev)
    torch.cuda.synchronize()
    p1, p2 = fe.pair_points(1)
    E, R, tv, ninl, good, status = fe.pair_pose(1)
    Ecv, mcv = cv2.findEssentialMat(p1, p2, K, cv2.RANSAC, 0.999, 1.0, 1000)
    g, _, _, _ = cv2.recoverPose(Ecv, p1, p2, K)
    assert abs(ninl - int(mcv.sum())) <= max(2, len(p1) // 100) and (status == 3) == (g < 8)
