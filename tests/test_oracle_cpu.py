"""CPU: the oracle restatement against the committed golden vectors (outputs of the reference's own
OpenCV code, tests/golden/make_golden.py) and — where cv2 is importable — against cv2 live."""
import glob
import os

import numpy as np
import pytest

from oracle import restated as R, synth, cv_ref, c_oracle
from tests.util import compare_orb

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_orb(path):
    g = np.load(path)
    k = g["kps"]
    ref = {"x": k[:, 0].astype(np.float32), "y": k[:, 1].astype(np.float32), "size": k[:, 2].astype(np.float32),
           "angle": k[:, 3].astype(np.float32), "response": k[:, 4].astype(np.float32),
           "octave": k[:, 5].astype(np.int32), "class_id": k[:, 6].astype(np.int32), "desc": g["desc"]}
    return g, ref


def as_struct(o):
    k = np.zeros(len(o["x"]), dtype=[("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                                      ("response", "<f4"), ("octave", "<i4")])
    for f in k.dtype.names:
        k[f] = o[f]
    return k


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "orb_*.npz"))))
def test_restated_orb_matches_golden(path):
    g, ref = load_orb(path)
    h, w, nf = int(g["h"]), int(g["w"]), int(g["nfeatures"])
    img = g["image"]
    assert np.array_equal(img, synth.frame(h, w, int(g["seed"]), bool(g["sparse"]))), "generator drifted"
    o = R.orb_detect_and_compute(img, nf, return_stages=True)
    st = compare_orb(as_struct(o), o["desc"], ref, h, w)
    assert st["missing"] == 0 and st["extra"] == 0, st
    if st["common"]:
        assert st["resp_eq"] == st["common"] and st["angle_eq"] == st["common"], st
        assert st["pt_eq"] == st["common"] and st["size_eq"] == st["common"], st
        assert st["desc_bits_diff"] == 0, st
    assert np.all(ref["class_id"] == -1)
    # stage hooks
    assert np.array_equal(o["stages"]["levels"][1], g["level1"])
    xs, ys, sc = o["stages"]["cand"][0]
    mine = set(zip(xs.tolist(), ys.tolist(), sc.tolist()))
    theirs = set(map(tuple, g["fast0"].tolist()))
    assert mine == theirs
    assert np.array_equal(o["stages"]["blurred"][0], g["blur0"])


def test_restated_match_matches_golden():
    g = np.load(os.path.join(GOLD, "match_700x900.npz"))
    q, t = g["q"], g["t"]
    m1 = R.match_k1(q, t)
    assert np.array_equal(m1[:, 0], g["match"][:, 0].astype(np.int32))
    assert np.array_equal(m1[:, 1], g["match"][:, 1].astype(np.int32))
    assert np.array_equal(m1[:, 2], g["match"][:, 3].astype(np.int32))
    assert np.all(g["match"][:, 2] == 0)
    idx, dist, cnt = R.knn(q, t, 2)
    assert np.array_equal(idx, g["knn2"][:, :, 1].astype(np.int32))
    assert np.array_equal(dist, g["knn2"][:, :, 3].astype(np.int32))
    x = R.cross_check(q, t)
    assert np.array_equal(x[:, :2], g["cross"][:, :2].astype(np.int32))
    # the plain-C restatement agrees with the numpy one (ties included)
    ci, cd, cc = c_oracle.knn(q, t, 2)
    assert np.array_equal(ci, idx) and np.array_equal(cd, dist) and np.array_equal(cc, cnt)


def test_quotas_and_level_sizes():
    # SURVEY.md Appendix A S0 tables
    assert R.quotas(500) == [109, 90, 75, 63, 52, 44, 36, 31]
    assert R.quotas(2000) == [434, 362, 302, 251, 209, 175, 145, 122]
    assert R.quotas(5000) == [1086, 905, 754, 628, 524, 436, 364, 303]
    assert [(w, h) for h, w, _ in R.level_sizes(376, 1241)] == [
        (1241, 376), (1034, 313), (862, 261), (718, 218), (598, 181), (499, 151), (416, 126), (346, 105)]
    assert [(w, h) for h, w, _ in R.level_sizes(1080, 1920)] == [
        (1920, 1080), (1600, 900), (1333, 750), (1111, 625), (926, 521), (772, 434), (643, 362), (536, 301)]


def test_reference_invariants_on_oracle():
    # /root/reference/src/feature/detector.rs:58-80 and matcher.rs:79-86
    blank = np.zeros((480, 640), np.uint8)
    o = R.orb_detect_and_compute(blank, 500)
    assert len(o["x"]) == 0 and o["desc"].shape == (0, 32)
    e = np.zeros((0, 32), np.uint8)
    assert len(R.match_k1(e, e)) == 0


def test_filter_good_matches_semantics():
    # matcher.rs:40-66: thr = max(ratio * min_dist, 30), strict <
    m = np.array([[0, 1, 10], [1, 2, 19], [2, 0, 20], [3, 5, 29], [4, 4, 30], [5, 3, 45]], np.int32)
    out = R.filter_good_matches(m, 2.0)   # thr = max(20, 30) = 30
    assert out[:, 0].tolist() == [0, 1, 2, 3]
    m2 = m.copy(); m2[:, 2] += 10         # min 20 -> thr 40
    assert R.filter_good_matches(m2, 2.0)[:, 0].tolist() == [0, 1, 2, 3]
    assert len(R.filter_good_matches(m[:0], 2.0)) == 0


def test_ratio_test_semantics():
    idx = np.array([[1, 2], [3, 4], [5, -1]], np.int32)
    dist = np.array([[30, 40], [30, 41], [3, -1]], np.int32)
    cnt = np.array([2, 2, 1], np.int32)
    keep = R.ratio_test(idx, dist, cnt, 0.75)   # 30 < 30.0 false ; 30 < 30.75 true ; single neighbour dropped
    assert keep.tolist() == [False, True, False]


@pytest.mark.skipif(not cv_ref.HAVE_CV2, reason="cv2 not importable")
@pytest.mark.parametrize("h,w,seed,nf,sparse", [(376, 1241, 0, 2000, False), (240, 320, 21, 500, False),
                                                (300, 400, 22, 1000, True)])
def test_restated_vs_cv2_live(h, w, seed, nf, sparse):
    img = synth.frame(h, w, seed, sparse)
    o = R.orb_detect_and_compute(img, nf)
    c = cv_ref.orb_detect_and_compute(img, nf)
    st = compare_orb(as_struct(o), o["desc"], c, h, w)
    assert st["missing"] == 0 and st["extra"] == 0, st
    assert st["resp_eq"] == st["common"] and st["angle_eq"] == st["common"] and st["desc_bits_diff"] == 0, st


@pytest.mark.skipif(not cv_ref.HAVE_CV2, reason="cv2 not importable")
def test_restated_match_vs_cv2_live():
    m, f = synth.map_and_frame_descriptors(3000, 400, f=1)
    assert np.array_equal(R.match_k1(m, f), cv_ref.match_k1(m, f))          # reference orientation (map.rs:107)
    i1, d1, c1 = R.knn(f, m, 2)
    i2, d2, c2 = cv_ref.knn2(f, m)
    assert np.array_equal(i1, i2) and np.array_equal(d1, d2) and np.array_equal(c1, c2)
    assert np.array_equal(R.cross_check(f, m), cv_ref.match_k1(f, m, cross_check=True))


def bgr_frame(h, w, seed):
    """Three differently seeded textures as B, G, R planes (deterministic)."""
    return np.stack([synth.frame(h, w, seed + 100 * c) for c in range(3)], axis=-1)


@pytest.mark.skipif(not cv_ref.HAVE_CV2, reason="cv2 not importable")
def test_bgr_to_gray_vs_cv2_live():
    import cv2
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (97, 131, 3), dtype=np.uint8)
    assert np.array_equal(R.bgr_to_gray(img), cv2.cvtColor(img, cv2.COLOR_BGR2GRAY))
    v = np.arange(0, 256, 5, dtype=np.uint8)                       # a grid over the whole colour cube
    cube = np.stack(np.meshgrid(v, v, v, indexing="ij"), -1).reshape(-1, 1, 3)
    assert np.array_equal(R.bgr_to_gray(cube), cv2.cvtColor(np.ascontiguousarray(cube), cv2.COLOR_BGR2GRAY))
    # cv::ORB on a BGR image == cv::ORB on its gray conversion (the implicit conversion the reference relies on)
    bgr = bgr_frame(200, 260, 40)
    orb = cv2.ORB_create(nfeatures=300)
    k1, d1 = orb.detectAndCompute(bgr, None)
    k2, d2 = orb.detectAndCompute(R.bgr_to_gray(bgr), None)
    assert [k.pt for k in k1] == [k.pt for k in k2] and np.array_equal(d1, d2)
