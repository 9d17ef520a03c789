"""Generates tests/golden/*.npz from the reference's own arithmetic (cv2 = the OpenCV C++ that
/root/reference/src/feature/{detector,matcher}.rs bind through the `opencv` crate).

Run in the build container (cv2 4.13.0):  python tests/golden/make_golden.py
The fixtures pin the oracle restatement (tests/test_oracle_cpu.py) and the CUDA path (tests/test_gpu_*.py)
wherever cv2 is not importable; where it is, the same comparisons also run live.
"""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def orb_case(name, h, w, seed, nfeatures, sparse=False):
    img = synth.frame(h, w, seed, sparse)
    orb = cv2.ORB_create(nfeatures=nfeatures)
    kps, desc = orb.detectAndCompute(img, None)
    kd = orb.detect(img, None)
    assert [k.pt for k in kd] == [k.pt for k in kps]
    arr = np.array([(k.pt[0], k.pt[1], k.size, k.angle, k.response, k.octave, k.class_id) for k in kps],
                   dtype=np.float64).reshape(-1, 7)
    if desc is None:
        desc = np.zeros((0, 32), np.uint8)
    # stage hooks: resize chain, FAST level 0, blurred level 0 (sepFilter2D fp32 path == what ORB runs)
    lv1 = cv2.resize(img, (int(round(w / 1.2000000476837158)), int(round(h / 1.2000000476837158))),
                     interpolation=cv2.INTER_LINEAR_EXACT)
    fast = cv2.FastFeatureDetector_create(20, True).detect(img, None)
    fa = np.array([(k.pt[0], k.pt[1], k.response) for k in fast], np.int32).reshape(-1, 3)
    k7 = cv2.getGaussianKernel(7, 2, cv2.CV_32F)
    blur0 = cv2.sepFilter2D(img, cv2.CV_8U, k7, k7, borderType=cv2.BORDER_REFLECT_101)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), h=h, w=w, seed=seed, nfeatures=nfeatures, sparse=sparse,
                        image=img, kps=arr, desc=desc, level1=lv1, fast0=fa, blur0=blur0,
                        cv2_version=cv2.__version__)
    print(name, "keypoints", len(kps), "fast0", len(fa))


def match_case(name, nq, nt, seed):
    rng = np.random.default_rng(seed)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    # forced ties / duplicates
    t[nt // 2] = t[nt - 3] = t[10]
    q[5] = t[10]
    q[7] = q[8]
    # make many near neighbours so the distance distribution is not degenerate
    for i in range(0, nq, 3):
        j = int(rng.integers(0, nt))
        flips = rng.random(256) < 0.08
        q[i] = t[j] ^ np.packbits(flips, bitorder="little")
    bf = cv2.BFMatcher(cv2.NORM_HAMMING, False)
    m1 = np.array([(d.queryIdx, d.trainIdx, d.imgIdx, d.distance) for d in bf.match(q, t)], np.float64)
    knn = bf.knnMatch(q, t, k=2)
    k2 = np.array([[(d.queryIdx, d.trainIdx, d.imgIdx, d.distance) for d in row] for row in knn], np.float64)
    mx = np.array([(d.queryIdx, d.trainIdx, d.imgIdx, d.distance)
                   for d in cv2.BFMatcher(cv2.NORM_HAMMING, True).match(q, t)], np.float64)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), q=q, t=t, match=m1, knn2=k2, cross=mx,
                        cv2_version=cv2.__version__)
    print(name, "k1", m1.shape, "knn2", k2.shape, "cross", mx.shape)


if __name__ == "__main__":
    orb_case("orb_240x320_s5_n300", 240, 320, 5, 300)
    orb_case("orb_200x260_s9_n500_sparse", 200, 260, 9, 500, sparse=True)
    orb_case("orb_70x200_s3_n100", 70, 200, 3, 100)
    match_case("match_700x900", 700, 900, 11)
    match_case("match_33x1", 33, 1, 12) if False else None
