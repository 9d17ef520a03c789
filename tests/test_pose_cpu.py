"""CPU: the oracle's restatement of the OpenCV calls behind odometry::PoseEstimator (oracle/pose_restated.py,
SURVEY.md §8 f2; /root/reference/src/odometry/pose.rs:47-114) pinned against cv2 itself: five-point solutions,
findEssentialMat (same RANSAC draws, same winner, same mask) and recoverPose."""
import numpy as np
import pytest

from oracle import pose_restated as PR

cv2 = pytest.importorskip("cv2")
KITTI = (718.856, 718.856, 607.1928, 185.2157)
K = np.array([[KITTI[0], 0, KITTI[2]], [0, KITTI[1], KITTI[3]], [0, 0, 1.0]])


def two_view_scene(n, seed, outliers=0.25, noise=0.3):
    rng = np.random.default_rng(seed)
    R = cv2.Rodrigues(rng.normal(0, 0.02, 3))[0]
    t = rng.normal(0, 0.3, 3) + [0, 0, 1.0]
    X = np.stack([rng.uniform(-10, 10, n), rng.uniform(-3, 3, n), rng.uniform(4, 40, n)], 1)

    def proj(Rm, tv):
        c = X @ Rm.T + tv
        return c[:, :2] / c[:, 2:] * [K[0, 0], K[1, 1]] + [K[0, 2], K[1, 2]]
    p1 = (proj(np.eye(3), np.zeros(3)) + rng.normal(0, noise, (n, 2))).astype(np.float32)
    p2 = (proj(R, t) + rng.normal(0, noise, (n, 2))).astype(np.float32)
    out = rng.random(n) < outliers
    p2[out] += rng.normal(0, 30, (int(out.sum()), 2)).astype(np.float32)
    return p1, p2, R, t


def test_five_point_vs_cv2():
    rng = np.random.default_rng(0)
    close, total = 0, 0
    for _ in range(60):
        R = cv2.Rodrigues(rng.normal(0, 0.2, 3))[0]
        t = rng.normal(0, 1, 3)
        X = np.stack([rng.uniform(-3, 3, 5), rng.uniform(-2, 2, 5), rng.uniform(3, 12, 5)], 1)
        x1 = X[:, :2] / X[:, 2:]
        c = X @ R.T + t
        x2 = c[:, :2] / c[:, 2:]
        sols = PR.five_point(x1, x2)
        want = cv2.findEssentialMat(x1.copy(), x2.copy(), np.eye(3), cv2.RANSAC, 0.999, 1.0)[0].reshape(-1, 3, 3)
        assert len(sols) == len(want)                                          # same number of real solutions
        for E in want:
            d = min(min(np.abs(E - S).max(), np.abs(E + S).max()) for S in sols)
            close += d < 1e-6
            total += 1
    assert close >= 0.97 * total


def test_cv_rng_samples_are_distinct_and_deterministic():
    s = PR.ransac_samples(37, 200)
    assert s.shape == (200, 5) and s.min() >= 0 and s.max() < 37
    assert all(len(set(r)) == 5 for r in s.tolist())
    assert np.array_equal(s, PR.ransac_samples(37, 200))


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_find_essential_mat_vs_cv2(seed):
    p1, p2, _, _ = two_view_scene(500, seed)
    Ecv, mcv = cv2.findEssentialMat(p1, p2, K, cv2.RANSAC, 0.999, 1.0, 1000)
    E, m, iters = PR.find_essential_mat(p1, p2, KITTI)
    assert iters < 1000
    assert min(np.abs(E - Ecv).max(), np.abs(E + Ecv).max()) < 1e-6            # the same hypothesis wins
    assert np.array_equal(m, mcv.ravel().astype(bool))


@pytest.mark.parametrize("seed", [4, 5])
def test_recover_pose_vs_cv2(seed):
    p1, p2, Rt, tt = two_view_scene(400, seed, outliers=0.1)
    Ecv, _ = cv2.findEssentialMat(p1, p2, K, cv2.RANSAC, 0.999, 1.0, 1000)
    n, Rcv, tcv, mcv = cv2.recoverPose(Ecv, p1, p2, K)
    R, t, m, good = PR.recover_pose(Ecv, p1, p2, KITTI)
    assert good == n and np.array_equal(m, mcv.ravel() != 0)
    assert np.abs(R - Rcv).max() < 1e-9 and np.abs(t - tcv.ravel()).max() < 1e-9
    assert np.abs(R - Rt).max() < 0.02                                         # and it is the true motion


def golden_pose_cases():
    """tests/golden/pose_kitti.npz (made by tests/golden/make_golden_pose.py with cv2 4.13.0): yields dicts of one case."""
    import os
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pose_kitti.npz"))
    for i in range(len(g["cases"])):
        yield {k[:-len(f"_{i}")]: g[k] for k in g.files if k.endswith(f"_{i}")}


def test_oracle_vs_committed_golden():
    """The restatement against the committed OpenCV outputs (no cv2 call here): same winner, mask, pose."""
    n = 0
    for c in golden_pose_cases():
        E, m, iters = PR.find_essential_mat(c["p1"], c["p2"], KITTI)
        assert min(np.abs(E - c["E"]).max(), np.abs(E + c["E"]).max()) < 1e-6
        assert np.array_equal(m, c["mask"] != 0) and 0 < iters <= 1000
        R, t, pm, good = PR.recover_pose(c["E"], c["p1"], c["p2"], KITTI)
        assert good == int(c["good"]) and np.array_equal(pm, c["pose_mask"] != 0)
        assert np.abs(R - c["R"]).max() < 1e-9 and np.abs(t - c["t"]).max() < 1e-9
        n += 1
    assert n == 3


def _host_solver(extra_flags=()):
    """The five-point solver and root finder of csrc/pose.cu (the section between the ORBX_HOST_TESTABLE markers is
    plain C++) compiled for the host — a test harness for the device code's arithmetic, not a product path."""
    import ctypes as C
    import os
    import subprocess
    import tempfile
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    s = open(os.path.join(root, "slam-rs_b200", "csrc", "pose.cu")).read()
    tables = s[s.index("// ORBX_HOST_TESTABLE_BEGIN"):s.index("// ORBX_HOST_TESTABLE_SPLIT")]
    body = s[s.index("// ORBX_HOST_TESTABLE_RESUME"):s.index("// ORBX_HOST_TESTABLE_END")]
    src = ("#include <math.h>\n#include <algorithm>\n#define __device__\n#define __constant__ static const\n"
           "#define __forceinline__ inline\n#define __restrict__\nconstexpr int kPoseMaxModels = 10;\n" + tables + body +
           "extern \"C\" int solve5(const double* x1, const double* x2, double* m) { double a[5][2], b[5][2];\n"
           "  for (int i = 0; i < 5; ++i) { a[i][0] = x1[2*i]; a[i][1] = x1[2*i+1]; b[i][0] = x2[2*i]; b[i][1] = x2[2*i+1]; }\n"
           "  double sm[kSolveScratch]; return five_point_solve<1>(a, b, m, sm); }\n"
           "extern \"C\" int roots(const double* c, int deg, double* r) { double d[121]; return real_roots<1>(c, deg, r, d); }\n")
    d = tempfile.mkdtemp(prefix="orbx_pose_host_")
    open(os.path.join(d, "host.cpp"), "w").write(src)
    subprocess.check_call(["g++", "-O2", "-shared", "-fPIC", *extra_flags, "-o", os.path.join(d, "libhost.so"),
                           os.path.join(d, "host.cpp")])
    return C.CDLL(os.path.join(d, "libhost.so"))


def test_device_solver_arithmetic_on_host():
    import ctypes as C
    lib = _host_solver()
    rng = np.random.default_rng(0)
    for _ in range(200):                                                       # real roots of random degree-10 polynomials
        c = rng.normal(0, 1, 11)
        r = np.zeros(10)
        n = lib.roots(C.c_void_p(c.ctypes.data), 10, C.c_void_p(r.ctypes.data))
        ref = np.roots(c[::-1])
        ref = np.sort(ref[np.abs(ref.imag) < 1e-10].real)
        assert n == len(ref) and (n == 0 or np.abs((r[:n] - ref) / np.maximum(1, np.abs(ref))).max() < 1e-8)

    def solve(x1, x2):
        out = np.zeros((10, 9))
        n = lib.solve5(C.c_void_p(np.ascontiguousarray(x1).ctypes.data), C.c_void_p(np.ascontiguousarray(x2).ctypes.data),
                       C.c_void_p(out.ctypes.data))
        return [out[i].reshape(3, 3) for i in range(n)]
    for seed in (11, 12):                                                      # the whole RANSAC with this solver == cv2
        p1, p2, _, _ = two_view_scene(500, seed)
        Ecv, mcv = cv2.findEssentialMat(p1, p2, K, cv2.RANSAC, 0.999, 1.0, 1000)
        E, m, _ = PR.find_essential_mat(p1, p2, KITTI, solver=solve)
        assert min(np.abs(E - Ecv).max(), np.abs(E + Ecv).max()) < 1e-6
        assert np.array_equal(m, mcv.ravel().astype(bool))


def test_interleaved_bisection_is_bit_identical_to_serial():
    """The root finder bisects four brackets of a level at a time (ORBX_ROOT_LANES = 4, for instruction-level parallelism
    on the device); with one lane it is the plain one-bracket-at-a-time bisection.  Same roots and same essential
    matrices, bit for bit - incl. polynomials with exact zeros at bracket ends and multiple roots."""
    import ctypes as C
    par, ser = _host_solver(), _host_solver(("-DORBX_ROOT_LANES=1",))
    rng = np.random.default_rng(5)
    for t in range(4000):
        deg = int(rng.integers(2, 11))
        c = np.zeros(11)
        c[:deg + 1] = np.poly(rng.integers(-3, 4, deg))[::-1] if t % 5 == 0 else rng.normal(0, 1, deg + 1)
        r1, r2 = np.zeros(10), np.zeros(10)
        n1 = ser.roots(C.c_void_p(c.ctypes.data), 10, C.c_void_p(r1.ctypes.data))
        n2 = par.roots(C.c_void_p(c.ctypes.data), 10, C.c_void_p(r2.ctypes.data))
        assert n1 == n2 and np.array_equal(r1[:n1].view(np.int64), r2[:n2].view(np.int64)), c
    for _ in range(500):
        x1 = rng.normal(0, 0.3, (5, 2))
        x2 = x1 + rng.normal(0, 0.05, (5, 2))
        m1, m2 = np.zeros((10, 9)), np.zeros((10, 9))
        n1 = ser.solve5(C.c_void_p(x1.ctypes.data), C.c_void_p(x2.ctypes.data), C.c_void_p(m1.ctypes.data))
        n2 = par.solve5(C.c_void_p(x1.ctypes.data), C.c_void_p(x2.ctypes.data), C.c_void_p(m2.ctypes.data))
        assert n1 == n2 and np.array_equal(m1[:n1].view(np.int64), m2[:n2].view(np.int64))


def test_cheirality_sign_symmetry_of_the_jacobi_null_vector():
    """pose_cheirality_kernel runs one SVD for the candidates (R, t) and (R, -t): their DLT systems differ by the sign
    of column 3, and csrc/jacobi.cuh's iteration is sign-symmetric in IEEE arithmetic, so the null vector of the second
    is (q0, q1, q2, -q3) or its negative BIT FOR BIT.  Checked on the host build of jacobi.cuh, with and without FMA
    contraction (the device contracts)."""
    import ctypes as C
    import os
    import subprocess
    import tempfile
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = ("#include <math.h>\n#define __device__\n#define __forceinline__ inline\n"
           "#include \"%s\"\n"
           "extern \"C\" void nullvec(const double* a, double* q) { double m[4][4], h[4];\n"
           "  for (int i = 0; i < 16; ++i) m[i / 4][i %% 4] = a[i];\n"
           "  orbx::null_vector4(m, h); for (int i = 0; i < 4; ++i) q[i] = h[i]; }\n"
           % os.path.join(root, "slam-rs_b200", "csrc", "jacobi.cuh"))
    d = tempfile.mkdtemp(prefix="orbx_jacobi_host_")
    open(os.path.join(d, "j.cpp"), "w").write(src)
    rng = np.random.default_rng(9)
    for flags in ((), ("-mfma", "-ffp-contract=fast")):
        so = os.path.join(d, "libj%d.so" % len(flags))
        subprocess.check_call(["g++", "-O2", "-shared", "-fPIC", *flags, "-o", so, os.path.join(d, "j.cpp")])
        lib = C.CDLL(so)
        for t in range(2000):
            if t % 2:                                                          # the DLT system of the kernel: [I | 0] and [R | t]
                ax, ay, bx, by = rng.normal(0, 0.4, 4)
                P = np.hstack([np.linalg.qr(rng.normal(0, 1, (3, 3)))[0], rng.normal(0, 1, (3, 1))])
                A = np.array([[-1, 0, ax, 0], [0, -1, ay, 0], bx * P[2] - P[0], by * P[2] - P[1]], np.float64)
            else:
                A = rng.normal(0, 1, (4, 4))
            A2 = A * np.array([1.0, 1.0, 1.0, -1.0])
            q, q2 = np.zeros(4), np.zeros(4)
            lib.nullvec(C.c_void_p(np.ascontiguousarray(A).ctypes.data), C.c_void_p(q.ctypes.data))
            lib.nullvec(C.c_void_p(np.ascontiguousarray(A2).ctypes.data), C.c_void_p(q2.ctypes.data))
            want = q * np.array([1.0, 1.0, 1.0, -1.0])
            assert (np.array_equal(q2.view(np.int64), want.view(np.int64))
                    or np.array_equal(q2.view(np.int64), (-want).view(np.int64))), (A, q, q2)
            sv = np.linalg.svd(A)
            if sv[1][2] - sv[1][3] > 1e-3 * sv[1][0]:                          # and it is LAPACK's smallest right singular vector
                v = sv[2][3]
                assert min(np.abs(q - v).max(), np.abs(q + v).max()) < 1e-9
