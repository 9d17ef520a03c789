"""TEST INFRASTRUCTURE ONLY — numpy restatement of the OpenCV calls behind odometry::PoseEstimator
(/root/reference/src/odometry/pose.rs:47-114): cv::findEssentialMat(RANSAC, 0.999, 1.0, 1000) and
cv::recoverPose.  The algorithm lives in OpenCV (calib3d/src/five-point.cpp, ptsetreg.cpp; `opencv` crate 0.97.1 ->
system OpenCV, pinned here against cv2 4.13.0 in tests/test_pose_cpu.py): its published structure is restated, not
copied —

  * points are normalised with K in f64 and the threshold divided by (fx + fy) / 2;
  * RANSAC draws 5 distinct indices per iteration from cv::RNG(state = 2^64 - 1) (multiply-with-carry, uniform(0, n) =
    next() % n, duplicates redrawn), evaluates every five-point solution on all points with the Sampson error rounded
    to fp32, keeps a model when its inlier count is strictly greater than the best so far (and > 4) and then shrinks
    the iteration budget with RANSACUpdateNumIters;
  * the five-point solver (Nister): null space of the 5 x 9 epipolar system, the ten cubic constraints
    det(E) = 0 and 2 E E^T E - tr(E E^T) E = 0 on E = x E1 + y E2 + z E3 + E4, Gauss-Jordan on the ten leading
    monomials, a 3 x 3 polynomial matrix B(z) with B(z) [x y 1]^T = 0, the real roots of det B(z) (degree 10);
  * recoverPose: E = U diag(s) V^T, R1 = U W V^T, R2 = U W^T V^T, t = u3, the four (R, +-t) candidates are scored by
    triangulating every point (cv::triangulatePoints, f64) with the depth tests of OpenCV (positive depth in both
    views, depth < 50), first maximum wins in the order (R1, t), (R2, t), (R1, -t), (R2, -t).
"""
import math

import numpy as np

LIN = [(1, 0, 0), (0, 1, 0), (0, 0, 1), (0, 0, 0)]
QUAD = [(2, 0, 0), (1, 1, 0), (1, 0, 1), (1, 0, 0), (0, 2, 0), (0, 1, 1), (0, 1, 0), (0, 0, 2), (0, 0, 1), (0, 0, 0)]
# column order of the 10 x 20 constraint matrix: x3 y3 x2y xy2 x2z x2 y2z y2 xyz xy | xz2 xz x yz2 yz y z3 z2 z 1
CUB = [(3, 0, 0), (0, 3, 0), (2, 1, 0), (1, 2, 0), (2, 0, 1), (2, 0, 0), (0, 2, 1), (0, 2, 0), (1, 1, 1), (1, 1, 0),
       (1, 0, 2), (1, 0, 1), (1, 0, 0), (0, 1, 2), (0, 1, 1), (0, 1, 0), (0, 0, 3), (0, 0, 2), (0, 0, 1), (0, 0, 0)]
_QI = {m: i for i, m in enumerate(QUAD)}
_CI = {m: i for i, m in enumerate(CUB)}
LL = np.array([[_QI[tuple(np.add(a, b))] for b in LIN] for a in LIN])     # linear x linear -> quadratic slot
QL = np.array([[_CI[tuple(np.add(a, b))] for b in LIN] for a in QUAD])    # quadratic x linear -> cubic slot


def _mul_ll(p, q):
    r = np.zeros(10)
    np.add.at(r, LL.ravel(), np.outer(p, q).ravel())
    return r


def _mul_ql(p, q):
    r = np.zeros(20)
    np.add.at(r, QL.ravel(), np.outer(p, q).ravel())
    return r


def constraint_matrix(Eb):
    """Eb: 4 x 9 null-space basis (rows E1..E4, row-major 3x3 each) -> the 10 x 20 coefficient matrix."""
    e = [[Eb[:, 3 * r + c] for c in range(3)] for r in range(3)]
    M = np.zeros((10, 20))

    def det2(a, b, c, d):
        return _mul_ll(a, d) - _mul_ll(b, c)
    M[0] = (_mul_ql(det2(e[1][1], e[1][2], e[2][1], e[2][2]), e[0][0])
            - _mul_ql(det2(e[1][0], e[1][2], e[2][0], e[2][2]), e[0][1])
            + _mul_ql(det2(e[1][0], e[1][1], e[2][0], e[2][1]), e[0][2]))
    EEt = [[sum(_mul_ll(e[i][k], e[j][k]) for k in range(3)) for j in range(3)] for i in range(3)]
    tr = EEt[0][0] + EEt[1][1] + EEt[2][2]
    L = [[EEt[i][j] - (0.5 * tr if i == j else 0.0) for j in range(3)] for i in range(3)]
    r = 1
    for i in range(3):
        for j in range(3):
            M[r] = sum(_mul_ql(L[i][k], e[k][j]) for k in range(3))
            r += 1
    return M


def five_point(x1, x2):
    """x1, x2: [5, 2] normalised image points -> list of 3x3 essential matrices of unit Frobenius norm."""
    Q = np.stack([x2[:, 0] * x1[:, 0], x2[:, 0] * x1[:, 1], x2[:, 0], x2[:, 1] * x1[:, 0], x2[:, 1] * x1[:, 1],
                  x2[:, 1], x1[:, 0], x1[:, 1], np.ones(5)], 1)
    Eb = np.linalg.svd(Q)[2][5:9]
    A = constraint_matrix(Eb)
    for c in range(10):
        p = c + int(np.argmax(np.abs(A[c:, c])))
        A[[c, p]] = A[[p, c]]
        if A[c, c] == 0.0:
            return []
        A[c] /= A[c, c]
        for rr in range(10):
            if rr != c:
                A[rr] -= A[rr, c] * A[c]
    R = A[:, 10:]

    def brow(e_, f_):  # (row of x^2 z-type) - z * (row of x^2-type); ascending powers of z
        return (np.array([e_[2], e_[1] - f_[2], e_[0] - f_[1], -f_[0]]),
                np.array([e_[5], e_[4] - f_[5], e_[3] - f_[4], -f_[3]]),
                np.array([e_[9], e_[8] - f_[9], e_[7] - f_[8], e_[6] - f_[7], -f_[6]]))
    B = [brow(R[4], R[5]), brow(R[6], R[7]), brow(R[8], R[9])]
    P = np.polynomial.polynomial
    pm = P.polymul
    det = (pm(B[0][0], P.polysub(pm(B[1][1], B[2][2]), pm(B[1][2], B[2][1])))
           - pm(B[0][1], P.polysub(pm(B[1][0], B[2][2]), pm(B[1][2], B[2][0])))
           + pm(B[0][2], P.polysub(pm(B[1][0], B[2][1]), pm(B[1][1], B[2][0]))))
    det = np.concatenate([det, np.zeros(11 - len(det))])
    sols = []
    for z in np.roots(det[::-1]):
        if abs(z.imag) > 1e-10:
            continue
        z = z.real
        Bz = np.array([[P.polyval(z, B[i][j]) for j in range(3)] for i in range(3)])
        v = max((np.cross(Bz[0], Bz[1]), np.cross(Bz[0], Bz[2]), np.cross(Bz[1], Bz[2])), key=lambda t: abs(t[2]))
        if abs(v[2]) < 1e-300:
            continue
        E = (v[0] / v[2]) * Eb[0] + (v[1] / v[2]) * Eb[1] + z * Eb[2] + Eb[3]
        sols.append((E / np.linalg.norm(E)).reshape(3, 3))
    return sols


class CvRNG:
    """cv::RNG: 64-bit multiply-with-carry, state = 2^64 - 1 in RANSACPointSetRegistrator::run."""

    def __init__(self, state=0xFFFFFFFFFFFFFFFF):
        self.state = state

    def next(self):
        self.state = ((self.state & 0xFFFFFFFF) * 4164903690 + (self.state >> 32)) & 0xFFFFFFFFFFFFFFFF
        return self.state & 0xFFFFFFFF

    def uniform(self, a, b):
        return a if a == b else int(self.next() % (b - a) + a)


def ransac_samples(n, iters, model_points=5):
    """The index subsets RANSAC draws for a set of n points: [iters, 5] (duplicates inside a subset are redrawn)."""
    rng = CvRNG()
    out = np.zeros((iters, model_points), np.int32)
    for it in range(iters):
        idx = []
        for _ in range(model_points):
            v = rng.uniform(0, n)
            while v in idx:
                v = rng.uniform(0, n)
            idx.append(v)
        out[it] = idx
    return out


def update_num_iters(p, ep, model_points, max_iters):
    p = min(max(p, 0.0), 1.0)
    ep = min(max(ep, 0.0), 1.0)
    num = max(1.0 - p, 2.2250738585072014e-308)
    denom = 1.0 - (1.0 - ep) ** model_points
    if denom < 2.2250738585072014e-308:
        return 0
    num, denom = math.log(num), math.log(denom)
    return max_iters if (denom >= 0 or -num >= max_iters * (-denom)) else int(np.rint(num / denom))


def normalize_points(pts, intr):
    fx, fy, cx, cy = intr
    p = np.asarray(pts, np.float32).reshape(-1, 2).astype(np.float64)
    return np.stack([(p[:, 0] - cx) / fx, (p[:, 1] - cy) / fy], 1)


def sampson_errors(E, x1, x2):
    h1 = np.hstack([x1, np.ones((len(x1), 1))])
    h2 = np.hstack([x2, np.ones((len(x2), 1))])
    Ex1 = h1 @ E.T
    Etx2 = h2 @ E
    num = (h2 * Ex1).sum(1)
    return (num * num / (Ex1[:, 0] ** 2 + Ex1[:, 1] ** 2 + Etx2[:, 0] ** 2 + Etx2[:, 1] ** 2)).astype(np.float32)


def find_essential_mat(pts1, pts2, intr, prob=0.999, threshold=1.0, max_iters=1000, solver=five_point):
    """cv::findEssentialMat(points1, points2, K, RANSAC, prob, threshold, maxIters, mask) for n > 5 points.
    Returns (E 3x3 or None, mask bool [n], iterations run)."""
    x1, x2 = normalize_points(pts1, intr), normalize_points(pts2, intr)
    n = len(x1)
    thr = threshold / ((intr[0] + intr[1]) / 2.0)
    t2 = thr * thr
    samples = ransac_samples(n, max_iters)
    niters, best, best_mask, max_good, it = max_iters, None, np.zeros(n, bool), 0, 0
    while it < niters:
        idx = samples[it]
        for E in solver(x1[idx], x2[idx]):
            m = sampson_errors(E, x1, x2).astype(np.float64) <= t2
            g = int(m.sum())
            if g > max(max_good, 4):
                best, best_mask, max_good = E, m, g
                niters = update_num_iters(prob, (n - g) / n, 5, niters)
        it += 1
    return best, best_mask, it


def _triangulate_f64(P0, P1, x1, x2):
    A = np.stack([x1[:, 0:1] * P0[2] - P0[0], x1[:, 1:2] * P0[2] - P0[1],
                  x2[:, 0:1] * P1[2] - P1[0], x2[:, 1:2] * P1[2] - P1[1]], 1)
    return np.linalg.svd(A)[2][:, 3, :]


def decompose_essential(E):
    U, _, Vt = np.linalg.svd(np.asarray(E, np.float64).reshape(3, 3))
    if np.linalg.det(U) < 0:
        U = -U
    if np.linalg.det(Vt) < 0:
        Vt = -Vt
    W = np.array([[0.0, 1, 0], [-1, 0, 0], [0, 0, 1]])
    return U @ W @ Vt, U @ W.T @ Vt, U[:, 2].copy()


def recover_pose(E, pts1, pts2, intr, distance_thresh=50.0):
    """cv::recoverPose(E, points1, points2, K, R, t, mask) -> (R, t, mask bool [n], n_good)."""
    x1, x2 = normalize_points(pts1, intr), normalize_points(pts2, intr)
    R1, R2, t = decompose_essential(E)
    P0 = np.hstack([np.eye(3), np.zeros((3, 1))])
    cands = [(R1, t), (R2, t), (R1, -t), (R2, -t)]
    masks = []
    for R, tt in cands:
        P1 = np.hstack([R, tt[:, None]])
        Q = _triangulate_f64(P0, P1, x1, x2)
        m = Q[:, 2] * Q[:, 3] > 0
        with np.errstate(divide="ignore", invalid="ignore"):
            Qn = Q / Q[:, 3:4]
        m &= Qn[:, 2] < distance_thresh
        z2 = Qn @ P1[2]
        m &= (z2 > 0) & (z2 < distance_thresh)
        masks.append(m)
    good = [int(m.sum()) for m in masks]
    k = 0 if good[0] >= max(good[1:]) else 1 if good[1] >= max(good[2:]) and good[1] >= good[0] else 2 if good[2] >= good[3] and good[2] >= good[0] and good[2] >= good[1] else 3
    return cands[k][0], cands[k][1], masks[k], good[k]
