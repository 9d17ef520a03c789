"""First GPU parity slice: matcher and extractor against the oracle through the C ABI."""
import numpy as np
import pytest

from oracle import restated as R, synth
from tests.util import compare_orb

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pkg():
    import slam_rs_b200 as s
    return s


def test_match_k1_random(pkg):
    q, t = synth.random_descriptors(700, 1), synth.random_descriptors(900, 2)
    t[500] = t[700] = t[10]
    q[5] = t[10]
    m = pkg.FeatureMatcher()
    out = m.match_descriptors(q, t)
    ref = R.match_k1(q, t)
    assert len(out) == 700
    assert np.array_equal(out["query_idx"], ref[:, 0])
    assert np.array_equal(out["train_idx"], ref[:, 1])
    assert np.array_equal(out["distance"], ref[:, 2].astype(np.float32))
    assert np.all(out["img_idx"] == 0)


def test_knn2_random(pkg):
    q, t = synth.random_descriptors(1000, 3), synth.random_descriptors(3000, 4)
    m = pkg.FeatureMatcher()
    out, cnt = m.knn_match(q, t)
    idx, dist, c = R.knn(q, t, 2)
    assert np.array_equal(cnt, c)
    assert np.array_equal(out["train_idx"], idx)
    assert np.array_equal(out["distance"], dist.astype(np.float32))


def test_orb_small(pkg):
    img = synth.frame(240, 320, 5)
    det = pkg.OrbDetector(300, max_width=320, max_height=240)
    kps, desc = det.detect_and_compute(img)
    ref = R.orb_detect_and_compute(img, 300)
    st = compare_orb(kps, desc, ref, 240, 320)
    print(st)
    assert st["missing"] == 0 and st["extra"] == 0
    assert st["resp_eq"] == st["common"] and st["angle_eq"] == st["common"]
    assert st["desc_bits_diff"] == 0
