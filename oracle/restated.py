"""numpy restatement of the OpenCV arithmetic behind the reference's ``feature`` module.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Each function follows one stage of
SURVEY.md Appendix A, i.e. what ``cv::ORB_Impl::detectAndCompute`` / ``cv::BFMatcher`` compute when
called from /root/reference/src/feature/detector.rs:20-40 and matcher.rs:24-66.  OpenCV itself is
an un-vendored system dependency of the reference (``opencv`` crate 0.97.1, Cargo.lock:5585-5588);
the restatement is pinned against cv2 4.13.0 in tests/test_oracle_cpu.py and tests/golden/.
"""
import os
import numpy as np

f32 = np.float32

NLEVELS = 8
SCALE_FACTOR = 1.2
EDGE = 31
FAST_T = 20
HARRIS_K = f32(0.04)

_here = os.path.dirname(os.path.abspath(__file__))
PATTERN = np.loadtxt(os.path.join(_here, "brief_pattern.txt"), dtype=np.int64).reshape(256, 4)

RING = [(0, 3), (1, 3), (2, 2), (3, 1), (3, 0), (3, -1), (2, -2), (1, -3), (0, -3), (-1, -3), (-2, -2),
        (-3, -1), (-3, 0), (-3, 1), (-2, 2), (-1, 3)]
UMAX = [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
GAUSS7 = np.array([float.fromhex(s) for s in (
    "0x1.1f5f62p-4", "0x1.0c70fcp-3", "0x1.869472p-3", "0x1.ba95c0p-3",
    "0x1.869472p-3", "0x1.0c70fcp-3", "0x1.1f5f62p-4")], dtype=np.float32)


# ---------------------------------------------------------------- S0 geometry and quotas
def level_sizes(h, w, n=NLEVELS, sf=SCALE_FACTOR):
    """[(h_l, w_l, scale_l)] — cv::ORB_Impl level geometry (Appendix A S0)."""
    sf = f32(sf)
    out = []
    for l in range(n):
        scale = f32(np.power(np.float64(sf), np.float64(l)))
        inv = f32(1.0) / scale
        out.append((int(np.rint(f32(h) * inv)), int(np.rint(f32(w) * inv)), float(scale)))
    return out


def quotas(N, n=NLEVELS, sf=SCALE_FACTOR):
    """Per-level feature quotas N_l (Appendix A S0)."""
    sf = f32(sf)
    factor = f32(1.0 / np.float64(sf))
    nd = f32(N) * (f32(1) - factor) / (f32(1) - f32(np.power(np.float64(factor), np.float64(n))))
    q, s = [], 0
    for _ in range(n - 1):
        q.append(int(np.rint(nd)))
        s += q[-1]
        nd = f32(nd * factor)
    return q + [max(N - s, 0)]


# ---------------------------------------------------------------- S1 INTER_LINEAR_EXACT chain
def resize_coeffs(S, D):
    scale = 1.0 / (np.float64(D) / np.float64(S))
    ofs = np.zeros(D, np.int64)
    c1 = np.zeros(D, np.int64)
    for d in range(D):
        fv = scale * (d + 0.5) - 0.5
        iv = int(np.floor(fv))
        if iv >= 0 and S > 1:
            if iv < S - 1:
                ofs[d] = iv
                c1[d] = int(np.rint((fv - iv) * 256))
            else:
                ofs[d] = S - 1
    return ofs, c1


def resize_exact(img, dh, dw):
    sh, sw = img.shape
    ox, cx = resize_coeffs(sw, dw)
    oy, cy = resize_coeffs(sh, dh)
    I = img.astype(np.int64)
    H = I[:, ox] * (256 - cx) + I[:, np.minimum(ox + 1, sw - 1)] * cx
    V = H[oy] * (256 - cy)[:, None] + H[np.minimum(oy + 1, sh - 1)] * cy[:, None]
    return np.clip((V + 32768) >> 16, 0, 255).astype(np.uint8)


def pyramid(img, n=NLEVELS, sf=SCALE_FACTOR):
    sizes = level_sizes(img.shape[0], img.shape[1], n, sf)
    lv = [np.ascontiguousarray(img)]
    for l in range(1, n):
        lv.append(resize_exact(lv[-1], sizes[l][0], sizes[l][1]))
    return lv, sizes


# ---------------------------------------------------------------- S2 FAST-9/16 + NMS
def fast_score_map(img, t=FAST_T):
    """Score map (0 = not a corner), Appendix A S2."""
    h, w = img.shape
    S = np.zeros((h, w), np.int32)
    if h < 7 or w < 7:
        return S
    I = img.astype(np.int32)
    c = I[3:h - 3, 3:w - 3]
    d = np.stack([c - I[3 + dy:h - 3 + dy, 3 + dx:w - 3 + dx] for dx, dy in RING])
    d2 = np.concatenate([d, d[:8]])
    dark = np.full(c.shape, -999, np.int32)
    bright = np.full(c.shape, -999, np.int32)
    for s in range(16):
        arc = d2[s:s + 9]
        dark = np.maximum(dark, arc.min(0))
        bright = np.maximum(bright, (-arc).min(0))
    m = np.maximum(dark, bright)
    S[3:h - 3, 3:w - 3] = np.where(m > t, m - 1, 0)
    return S


def fast_nms(S):
    """(x, y, score) of strict 3x3 local maxima in raster order."""
    h, w = S.shape
    P = np.pad(S, 1)
    keep = S > 0
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            if dx == 0 and dy == 0:
                continue
            keep &= S > P[1 + dy:1 + dy + h, 1 + dx:1 + dx + w]
    ys, xs = np.nonzero(keep)
    return xs.astype(np.int64), ys.astype(np.int64), S[ys, xs].astype(np.int64)


# ---------------------------------------------------------------- S3/S5 retainBest
def retain_best_mask(vals, n):
    """cv::KeyPointsFilter::retainBest: keep all with value >= n-th largest (ties kept)."""
    vals = np.asarray(vals)
    if n < 0 or len(vals) <= n:
        return np.ones(len(vals), bool)
    if n == 0:
        return np.zeros(len(vals), bool)
    r = np.sort(vals)[::-1][n - 1]
    return vals >= r


# ---------------------------------------------------------------- S4 Harris
def harris_abc(img, xs, ys):
    I = img.astype(np.int64)
    a = np.zeros(len(xs), np.int64)
    b = np.zeros(len(xs), np.int64)
    c = np.zeros(len(xs), np.int64)
    for dy in range(-3, 4):
        for dx in range(-3, 4):
            X = xs + dx
            Y = ys + dy
            Ix = 2 * (I[Y, X + 1] - I[Y, X - 1]) + (I[Y - 1, X + 1] - I[Y - 1, X - 1]) + (I[Y + 1, X + 1] - I[Y + 1, X - 1])
            Iy = 2 * (I[Y + 1, X] - I[Y - 1, X]) + (I[Y + 1, X - 1] - I[Y - 1, X - 1]) + (I[Y + 1, X + 1] - I[Y - 1, X + 1])
            a += Ix * Ix
            b += Iy * Iy
            c += Ix * Iy
    return a, b, c


def harris_response(a, b, c, k=HARRIS_K):
    scale = f32(1.0) / (f32(4) * f32(7) * f32(255.0))
    s4 = f32(f32(f32(scale * scale) * scale) * scale)
    fa, fb, fc = np.asarray(a).astype(f32), np.asarray(b).astype(f32), np.asarray(c).astype(f32)
    t1 = (fa * fb).astype(f32) - (fc * fc).astype(f32)
    s = (fa + fb).astype(f32)
    t2 = ((k * s).astype(f32) * s).astype(f32)
    return ((t1.astype(f32) - t2).astype(f32) * s4).astype(f32)


# ---------------------------------------------------------------- S6 orientation
def ic_moments(img, xs, ys):
    I = img.astype(np.int64)
    m10 = np.zeros(len(xs), np.int64)
    m01 = np.zeros(len(xs), np.int64)
    for u in range(-15, 16):
        m10 += u * I[ys, xs + u]
    for v in range(1, 16):
        vs = np.zeros(len(xs), np.int64)
        for u in range(-UMAX[v], UMAX[v] + 1):
            p = I[ys + v, xs + u]
            q = I[ys - v, xs + u]
            m10 += u * (p + q)
            vs += p - q
        m01 += v * vs
    return m10, m01


_s = f32(180.0 / np.pi)
_P1 = f32(f32(0.9997878412794807) * _s)
_P3 = f32(f32(-0.3258083974640975) * _s)
_P5 = f32(f32(0.1555786518463281) * _s)
_P7 = f32(f32(-0.04432655554792128) * _s)
_EPS = f32(2.220446049250313e-16)


def fast_atan2(y, x):
    """cv::fastAtan2 (degrees) vectorised, non-fused fp32 (Appendix A S6)."""
    y = np.asarray(y).astype(f32)
    x = np.asarray(x).astype(f32)
    ax, ay = np.abs(x), np.abs(y)
    big = ax >= ay
    num = np.where(big, ay, ax).astype(f32)
    den = (np.where(big, ax, ay).astype(f32) + _EPS).astype(f32)
    c = (num / den).astype(f32)
    c2 = (c * c).astype(f32)
    a = (_P7 * c2).astype(f32)
    a = (a + _P5).astype(f32)
    a = (a * c2).astype(f32)
    a = (a + _P3).astype(f32)
    a = (a * c2).astype(f32)
    a = (a + _P1).astype(f32)
    a = (a * c).astype(f32)
    a = np.where(big, a, (f32(90.0) - a).astype(f32)).astype(f32)
    a = np.where(x < 0, (f32(180.0) - a).astype(f32), a).astype(f32)
    a = np.where(y < 0, (f32(360.0) - a).astype(f32), a).astype(f32)
    return a


# ---------------------------------------------------------------- S8 blur
def _fma(a, b, c):
    # single-rounding fused multiply-add emulated in float64 (a*b exact in f64 for f32 inputs)
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(f32)


def blur7(img):
    """sepFilter2D with the fp32 7-tap sigma=2 Gaussian, FMA accumulation, REFLECT_101."""
    k = GAUSS7
    E = np.pad(img, 3, mode="reflect").astype(f32)
    h, w = img.shape
    kk = [np.full((), kv, f32) for kv in k]
    acc = (kk[0] * E[:, 0:w]).astype(f32)
    for j in range(1, 7):
        acc = _fma(np.broadcast_to(kk[j], acc.shape), E[:, j:j + w], acc)
    R = acc  # (h+6, w)
    acc = (kk[3] * R[3:3 + h]).astype(f32)
    for j in range(1, 4):
        pair = (R[3 + j:3 + j + h] + R[3 - j:3 - j + h]).astype(f32)
        acc = _fma(np.broadcast_to(kk[3 + j], acc.shape), pair, acc)
    return np.clip(np.rint(acc), 0, 255).astype(np.uint8)


# ---------------------------------------------------------------- S9 rBRIEF
def describe(B, cx, cy, angle_deg, pat=PATTERN):
    """256-bit steered BRIEF descriptors for integer level coordinates (vectorised over keypoints)."""
    cx = np.asarray(cx, np.int64)
    cy = np.asarray(cy, np.int64)
    th = (np.asarray(angle_deg).astype(f32) * f32(np.pi / 180.0)).astype(f32)
    a = np.cos(th.astype(np.float64)).astype(f32)[:, None]
    b = np.sin(th.astype(np.float64)).astype(f32)[:, None]
    out = np.zeros((len(cx), 32), np.uint8)
    bits = []
    vals = []
    for col in (0, 2):
        px = pat[:, col].astype(f32)[None, :]
        py = pat[:, col + 1].astype(f32)[None, :]
        xr = ((px * a).astype(f32) - (py * b).astype(f32)).astype(f32)
        yr = ((px * b).astype(f32) + (py * a).astype(f32)).astype(f32)
        ix = np.rint(xr).astype(np.int64)
        iy = np.rint(yr).astype(np.int64)
        vals.append(B[cy[:, None] + iy, cx[:, None] + ix].astype(np.int64))
    bits = (vals[0] < vals[1]).astype(np.uint8)
    if len(cx):
        out = np.packbits(bits, axis=1, bitorder="little")
    return out


# ---------------------------------------------------------------- ingest: BGR -> gray in front of S1
def bgr_to_gray(img):
    """cv::cvtColor(COLOR_BGR2GRAY) for 8UC3, as called at /root/reference/examples/visual_odometry.rs:93-99 and
    src/main.rs:43-49 (and implicitly by cv::ORB for non-8UC1 input, src/feature/detector.rs:22,32).  15-bit fixed
    point: (B*3735 + G*19235 + R*9798 + 16384) >> 15.  Pinned against cv2 in tests/test_oracle_cpu.py."""
    a = np.asarray(img).astype(np.int64)
    return ((a[..., 0] * 3735 + a[..., 1] * 19235 + a[..., 2] * 9798 + 16384) >> 15).astype(np.uint8)


# ---------------------------------------------------------------- S1–S9 end to end
def orb_detect_and_compute(img, nfeatures=500, with_desc=True, return_stages=False):
    """Restated cv::ORB_Impl::detectAndCompute.  Returns a dict of arrays in level-major order;
    inside a level the order is (response desc, y, x) — the library's deterministic order."""
    lv, sizes = pyramid(img)
    q = quotas(nfeatures)
    rows = []
    stages = {"levels": lv, "sizes": sizes, "quotas": q, "cand": [], "sel1": []}
    for l, L in enumerate(lv):
        h, w = L.shape
        S = fast_score_map(L)
        xs, ys, sc = fast_nms(S)
        stages["cand"].append((xs, ys, sc))
        inb = (xs >= EDGE) & (xs < w - EDGE) & (ys >= EDGE) & (ys < h - EDGE)
        xs, ys, sc = xs[inb], ys[inb], sc[inb]
        m = retain_best_mask(sc, 2 * q[l])
        xs, ys, sc = xs[m], ys[m], sc[m]
        if len(xs) == 0:
            stages["sel1"].append((xs, ys, sc, np.zeros(0, f32)))
            continue
        a, b, c = harris_abc(L, xs, ys)
        resp = harris_response(a, b, c)
        stages["sel1"].append((xs, ys, sc, resp))
        m = retain_best_mask(resp, q[l])
        xs, ys, resp = xs[m], ys[m], resp[m]
        m10, m01 = ic_moments(L, xs, ys)
        ang = fast_atan2(m01, m10)
        order = np.lexsort((xs, ys, -resp.astype(np.float64)))
        for i in order:
            rows.append((l, int(xs[i]), int(ys[i]), resp[i], ang[i]))
    K = len(rows)
    out = {
        "octave": np.array([r[0] for r in rows], np.int32),
        "xl": np.array([r[1] for r in rows], np.int32),
        "yl": np.array([r[2] for r in rows], np.int32),
        "response": np.array([r[3] for r in rows], f32),
        "angle": np.array([r[4] for r in rows], f32),
    }
    sc = np.array([sizes[l][2] for l in out["octave"]], f32) if K else np.zeros(0, f32)
    out["x"] = (out["xl"].astype(f32) * sc).astype(f32)
    out["y"] = (out["yl"].astype(f32) * sc).astype(f32)
    out["size"] = (f32(31.0) * sc).astype(f32)
    if with_desc:
        desc = np.zeros((K, 32), np.uint8)
        blurred = [blur7(L) if (L.shape[0] > 6 and L.shape[1] > 6) else L for L in lv]
        stages["blurred"] = blurred
        for l in range(len(lv)):
            idx = np.nonzero(out["octave"] == l)[0]
            if len(idx):
                desc[idx] = describe(blurred[l], out["xl"][idx], out["yl"][idx], out["angle"][idx])
        out["desc"] = desc
    if return_stages:
        out["stages"] = stages
    return out


# ---------------------------------------------------------------- S10 matching
def hamming_matrix(q, t):
    q = np.ascontiguousarray(q, np.uint8)
    t = np.ascontiguousarray(t, np.uint8)
    lut = np.array([bin(i).count("1") for i in range(256)], np.int32)
    D = np.zeros((len(q), len(t)), np.int32)
    for b in range(32):
        D += lut[q[:, b][:, None] ^ t[:, b][None, :]]
    return D


def knn(q, t, k):
    """(idx[Q,k], dist[Q,k], count[Q]) — first k of a stable sort by distance (Appendix A S10)."""
    Q, T = len(q), len(t)
    kk = min(k, T)
    idx = np.full((Q, k), -1, np.int32)
    dist = np.full((Q, k), -1, np.int32)
    if Q == 0 or T == 0:
        return idx, dist, np.zeros(Q, np.int32)
    for s in range(0, Q, 1024):
        D = hamming_matrix(q[s:s + 1024], t)
        o = np.argsort(D, axis=1, kind="stable")[:, :kk]
        idx[s:s + 1024, :kk] = o
        dist[s:s + 1024, :kk] = np.take_along_axis(D, o, 1)
    return idx, dist, np.full(Q, kk, np.int32)


def match_k1(q, t):
    """FeatureMatcher::match_descriptors (matcher.rs:24-37): empty side -> no matches."""
    if len(q) == 0 or len(t) == 0:
        return np.zeros((0, 3), np.int32)
    idx, dist, _ = knn(q, t, 1)
    return np.stack([np.arange(len(q), dtype=np.int32), idx[:, 0], dist[:, 0]], 1)


def cross_check(q, t):
    """BFMatcher(crossCheck=true).match: mutual first-arg-min, ascending queryIdx."""
    if len(q) == 0 or len(t) == 0:
        return np.zeros((0, 3), np.int32)
    f = match_k1(q, t)
    r = match_k1(t, q)
    keep = r[f[:, 1], 1] == f[:, 0]
    return f[keep]


def ratio_test(idx, dist, count, ratio):
    """Lowe: keep query i iff it has two neighbours and d1 < ratio*d2 (fp32, strict)."""
    d1 = dist[:, 0].astype(f32)
    d2 = (f32(ratio) * dist[:, 1].astype(f32)).astype(f32)
    return (count >= 2) & (d1 < d2)


def filter_good_matches(matches, ratio):
    """FeatureMatcher::filter_good_matches (matcher.rs:40-66). matches: int array (n,3) q,t,dist."""
    if len(matches) == 0:
        return matches
    d = matches[:, 2].astype(f32)
    thr = max(f32(f32(ratio) * d.min()), f32(30.0))
    return matches[d < thr]


# ---------------------------------------------------------------- mapping::Map hot path
def map_project(points, R, t, fx, fy, cx, cy):
    """Map::project_point (/root/reference/src/mapping/map.rs:50-69) and the bounds test of find_matches (:90) for
    an [n, 3] float64 array.  nalgebra's r * point + t accumulates the columns left to right in f64 without fused
    multiply-adds: ((r_i0*X + r_i1*Y) + r_i2*Z) + t_i.  Returns (xy [n, 2], visible [n])."""
    P = np.asarray(points, np.float64).reshape(-1, 3)
    R = np.asarray(R, np.float64).reshape(3, 3)
    t = np.asarray(t, np.float64).reshape(3)
    cam = [((R[i, 0] * P[:, 0] + R[i, 1] * P[:, 1]) + R[i, 2] * P[:, 2]) + t[i] for i in range(3)]
    front = cam[2] > 0.0
    z = np.where(front, cam[2], 1.0)
    x = np.where(front, fx * (cam[0] / z) + cx, 0.0)
    y = np.where(front, fy * (cam[1] / z) + cy, 0.0)
    vis = front & (x >= 0.0) & (x < 4000.0) & (y >= 0.0) & (y < 3000.0)
    return np.stack([x, y], 1), vis


def map_find_matches(points, descs, has_desc, alive, R, t, intr, frame_desc):
    """Map::find_matches (map.rs:72-118): visible points with a descriptor, in ascending id; k=1 match with
    query = map rows, train = frame rows (:107); filter_good_matches(2.0) (:108); -> [(map_id, kp_idx)]."""
    _, vis = map_project(points, R, t, *intr)
    ids = np.nonzero(vis & np.asarray(has_desc, bool) & np.asarray(alive, bool))[0]
    if len(ids) == 0 or len(frame_desc) == 0:
        return np.zeros((0, 2), np.int64)
    good = filter_good_matches(match_k1(np.asarray(descs)[ids], frame_desc), 2.0)
    return np.stack([ids[good[:, 0]], good[:, 1]], 1) if len(good) else np.zeros((0, 2), np.int64)


# ---------------------------------------------------------------- §8 f2 (first step): matched points
def extract_matched_points(kp1_xy, kp2_xy, matches):
    """/root/reference/src/odometry/pose.rs:27-44: for every match (query_idx, train_idx) push kp1[query_idx].pt
    and kp2[train_idx].pt.  kp*_xy: float32 [n, 2]; matches: int [m, 2].  Raises IndexError like `Vector::get(..)?`."""
    p1, p2 = [], []
    for q, t in np.asarray(matches).reshape(-1, 2):
        if not (0 <= q < len(kp1_xy)) or not (0 <= t < len(kp2_xy)):
            raise IndexError("match index out of range")
        p1.append(kp1_xy[q])
        p2.append(kp2_xy[t])
    return (np.asarray(p1, np.float32).reshape(-1, 2), np.asarray(p2, np.float32).reshape(-1, 2))


# ---------------------------------------------------------------- §8 f4: mapping::Triangulator
def projection_matrix(intr, R, t):
    """Triangulator::build_projection_matrix (/root/reference/src/mapping/triangulation.rs:161-183): P = K [R | t]
    by cv::gemm (f64; small matrices take OpenCV's own loop: k ascending, no fused multiply-add).  K's zero
    entries contribute exact zeros, so row 0 = fx*Rt[0] + cx*Rt[2], row 1 = fy*Rt[1] + cy*Rt[2], row 2 = Rt[2]."""
    fx, fy, cx, cy = intr
    Rt = np.hstack([np.asarray(R, np.float64).reshape(3, 3), np.asarray(t, np.float64).reshape(3, 1)])
    return np.stack([fx * Rt[0] + cx * Rt[2], fy * Rt[1] + cy * Rt[2], Rt[2]])


def triangulate_points4d(P1, P2, pts1, pts2):
    """cv::triangulatePoints as the reference calls it (triangulation.rs:114): per correspondence the 4x4 system
    rows (x*P[2] - P[0], y*P[2] - P[1]) for both views in f64, SVD, last right singular vector, stored as CV_32F
    (the inputs are Point2f).  Returns float32 [n, 4]; the overall sign of each row is arbitrary."""
    p1 = np.asarray(pts1, np.float32).reshape(-1, 2).astype(np.float64)
    p2 = np.asarray(pts2, np.float32).reshape(-1, 2).astype(np.float64)
    A = np.stack([p1[:, 0:1] * P1[2] - P1[0], p1[:, 1:2] * P1[2] - P1[1],
                  p2[:, 0:1] * P2[2] - P2[0], p2[:, 1:2] * P2[2] - P2[1]], 1)
    if len(A) == 0:
        return np.zeros((0, 4), np.float32)
    return np.linalg.svd(A)[2][:, 3, :].astype(np.float32)


def triangulate(intr, pose1, pose2, pts1, pts2, descriptors=None):
    """Triangulator::triangulate (triangulation.rs:92-158).  Returns (ids, xyz float64 [m, 3], desc uint8 [m, 32],
    has_desc bool [m]): points with |w| >= 1e-10 (:127) in front of both cameras (:134-139, nalgebra's
    ((r20*x + r21*y) + r22*z) + t2 > 0), in input order with id = input index (:140) and descriptor row i attached
    when i < rows (:141-150).  Raises ValueError when the point arrays differ in length (:100-102)."""
    pts1 = np.asarray(pts1, np.float32).reshape(-1, 2)
    pts2 = np.asarray(pts2, np.float32).reshape(-1, 2)
    if len(pts1) != len(pts2):
        raise ValueError("Point arrays must have the same length")
    n = len(pts1)
    if n == 0:
        return np.zeros(0, np.int32), np.zeros((0, 3)), np.zeros((0, 32), np.uint8), np.zeros(0, bool)
    (R1, t1), (R2, t2) = [(np.asarray(r, np.float64).reshape(3, 3), np.asarray(t, np.float64).reshape(3)) for r, t in (pose1, pose2)]
    h = triangulate_points4d(projection_matrix(intr, R1, t1), projection_matrix(intr, R2, t2), pts1, pts2).astype(np.float64)
    w = h[:, 3]
    ok = ~(np.abs(w) < 1e-10)
    with np.errstate(divide="ignore", invalid="ignore"):
        X = h[:, :3] / w[:, None]
    for R, t in ((R1, t1), (R2, t2)):
        ok &= (((R[2, 0] * X[:, 0] + R[2, 1] * X[:, 1]) + R[2, 2] * X[:, 2]) + t[2]) > 0.0
    ids = np.nonzero(ok)[0].astype(np.int32)
    rows = 0 if descriptors is None else len(descriptors)
    desc = np.zeros((len(ids), 32), np.uint8)
    has = ids < rows
    if rows:
        desc[has] = np.asarray(descriptors, np.uint8).reshape(rows, 32)[ids[has]]
    return ids, X[ids], desc, has


def transform_points_to_world(xyz, pose):
    """transform_points_to_world (/root/reference/examples/point_cloud.rs:522-544): r_ctw = R^T, t_ctw = -r_ctw * t,
    p_world = r_ctw * p + t_ctw; nalgebra accumulates the columns left to right in f64, unfused."""
    R, t = np.asarray(pose[0], np.float64).reshape(3, 3), np.asarray(pose[1], np.float64).reshape(3)
    rc = R.T
    tc = ((-rc[:, 0]) * t[0] + (-rc[:, 1]) * t[1]) + (-rc[:, 2]) * t[2]
    p = np.asarray(xyz, np.float64).reshape(-1, 3)
    return ((rc[:, 0] * p[:, 0:1] + rc[:, 1] * p[:, 1:2]) + rc[:, 2] * p[:, 2:3]) + tc
