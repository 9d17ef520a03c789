"""The reference's own arithmetic through cv2 — TEST INFRASTRUCTURE / timed CPU baseline only.

``cv2.ORB_create(...).detectAndCompute`` and ``cv2.BFMatcher(NORM_HAMMING)`` execute the same
``cv::ORB_Impl`` / ``cv::BFMatcher`` code that the reference reaches through the ``opencv`` crate at
/root/reference/src/feature/detector.rs:16,22,32 and src/feature/matcher.rs:16,34.  The only
difference to the reference wrapper is that ``nfeatures`` is honoured (detector.rs:15-17 ignores
``max_features``) — ``nfeatures=500`` is the reference-faithful setting.
"""
import numpy as np

try:
    import cv2
    HAVE_CV2 = True
except Exception:  # pragma: no cover - cv2 is part of the image
    cv2 = None
    HAVE_CV2 = False


def set_threads(n):
    cv2.setNumThreads(int(n))


def orb_detect_and_compute(img, nfeatures=500):
    orb = cv2.ORB_create(nfeatures=nfeatures)
    kps, desc = orb.detectAndCompute(img, None)
    return _pack(kps, desc)


def orb_detect(img, nfeatures=500):
    orb = cv2.ORB_create(nfeatures=nfeatures)
    return _pack(orb.detect(img, None), None)


def _pack(kps, desc):
    K = len(kps)
    out = {
        "x": np.array([k.pt[0] for k in kps], np.float32),
        "y": np.array([k.pt[1] for k in kps], np.float32),
        "size": np.array([k.size for k in kps], np.float32),
        "angle": np.array([k.angle for k in kps], np.float32),
        "response": np.array([k.response for k in kps], np.float32),
        "octave": np.array([k.octave for k in kps], np.int32),
        "class_id": np.array([k.class_id for k in kps], np.int32),
    }
    if desc is None:
        desc = np.zeros((K, 32), np.uint8) if K == 0 else None
    out["desc"] = desc
    return out


def match_k1(q, t, cross_check=False):
    """[(queryIdx, trainIdx, distance)] int32 array; empty guard as matcher.rs:29-31."""
    if len(q) == 0 or len(t) == 0:
        return np.zeros((0, 3), np.int32)
    m = cv2.BFMatcher(cv2.NORM_HAMMING, cross_check).match(q, t)
    return np.array([(d.queryIdx, d.trainIdx, int(d.distance)) for d in m], np.int32).reshape(-1, 3)


def knn2(q, t):
    """idx[Q,2], dist[Q,2], count[Q] from BFMatcher.knnMatch(k=2).  Train is chunked below OpenCV's
    262144-row assert and merged on the (distance, trainIdx) key."""
    Q = len(q)
    idx = np.full((Q, 2), -1, np.int32)
    dist = np.full((Q, 2), -1, np.int32)
    cnt = np.zeros(Q, np.int32)
    if Q == 0 or len(t) == 0:
        return idx, dist, cnt
    CH = 262143
    cands = [[] for _ in range(Q)]
    bf = cv2.BFMatcher(cv2.NORM_HAMMING, False)
    for s in range(0, len(t), CH):
        res = bf.knnMatch(q, t[s:s + CH], k=2)
        for i, lst in enumerate(res):
            for d in lst:
                cands[i].append((int(d.distance), d.trainIdx + s))
    for i, c in enumerate(cands):
        c.sort()
        c = c[:2]
        cnt[i] = len(c)
        for j, (dd, ti) in enumerate(c):
            idx[i, j] = ti
            dist[i, j] = dd
    return idx, dist, cnt
