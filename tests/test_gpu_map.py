"""GPU: the device-resident map store (csrc/mapstore.cu) through the C ABI vs the oracle, plus
ports of the reference's own Map tests (/root/reference/src/mapping/map.rs:183-240)."""
import numpy as np
import pytest

from oracle import restated as R, synth

pytestmark = pytest.mark.gpu
KITTI = (718.856, 718.856, 607.1928, 185.2157)


@pytest.fixture(scope="module")
def orbx():
    import slam_rs_b200 as s
    return s


def rot(rx, ry, rz):
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def test_reference_map_tests(orbx):
    cam = orbx.CameraIntrinsics.kitti()
    m = orbx.Map(cam, capacity=16)
    assert m.size() == 0                                                     # test_map_creation (map.rs:186-190)
    ids = m.add_points([[1.0, 2.0, 10.0], [2.0, 3.0, 10.0]])                 # test_add_points (:193-204)
    assert m.size() == 2 and ids.tolist() == [0, 1]
    m2 = orbx.Map(cam, capacity=16)
    m2.add_points([[0.0, 0.0, 10.0]])
    xy, vis, alive, has = m2.project_points((np.eye(3), np.zeros(3)))        # test_projection (:207-221)
    assert vis[0] and alive[0] and not has[0]
    assert abs(xy[0, 0] - cam.cx) < 1e-6 and abs(xy[0, 1] - cam.cy) < 1e-6
    m3 = orbx.Map(cam, capacity=16)                                          # test_prune_outliers (:224-240)
    m3.add_points([[1.0, 2.0, 10.0], [2.0, 3.0, 10.0]], observations=[5, 1])
    assert m3.stable_count() == 1
    assert m3.prune_outliers() == 1 and m3.size() == 1
    m3.clear()
    assert m3.size() == 0 and m3.add_points([[0.0, 0.0, 1.0]]).tolist() == [0]
    with pytest.raises(orbx.OrbxError):                                      # capacity is a reported error
        orbx.Map(cam, capacity=2).add_points(np.zeros((3, 3)))


@pytest.mark.parametrize("n,nf,seed", [(20000, 700, 1), (3001, 129, 2)])
def test_find_matches_vs_oracle(orbx, n, nf, seed):
    rng = np.random.default_rng(seed)
    mp, fr = synth.map_and_frame_descriptors(n, nf, f=seed)
    pts = np.stack([rng.uniform(-40, 40, n), rng.uniform(-10, 10, n), rng.uniform(-20, 80, n)], 1)
    pts[:5] = [[0, 0, 0], [1, 1, -1e-300], [0, 0, 1e-300], [np.nan, 0, 1], [1e308, 0, 1]]   # edge inputs
    has = rng.random(n) < 0.93
    Rm, t = rot(0.02, -0.3, 0.01), np.array([0.3, -0.1, 1.5])
    m = orbx.Map(orbx.CameraIntrinsics(*KITTI), capacity=n + 8)
    half = n // 2
    m.add_points(pts[:half], mp[:half], has[:half])
    m.add_points(pts[half:], mp[half:], has[half:])
    xy, vis, alive, hd = m.project_points((Rm, t))
    rxy, rvis = R.map_project(pts, Rm, t, *KITTI)
    assert np.array_equal(vis, rvis) and np.array_equal(hd, has) and alive.all()
    assert np.array_equal(xy[rvis], rxy[rvis])                               # f64, bit exact
    matcher = orbx.FeatureMatcher()
    got = m.find_matches(None, fr, (Rm, t), matcher)
    want = R.map_find_matches(pts, mp, has, np.ones(n, bool), Rm, t, KITTI, fr)
    assert m.last_visible == int((rvis & has).sum()) and 0 < len(want) < m.last_visible
    assert np.array_equal(got, want)
    # observations / pruning follow map.rs:121-137
    m.update_observations(got)
    obs, alive = m.observations()
    exp = np.ones(n, np.int64)
    np.add.at(exp, want[:, 0], 1)
    assert np.array_equal(obs, exp)
    removed = m.prune_outliers()
    assert removed == int((exp < 2).sum()) and m.size() == n - removed
    # pruned points no longer take part
    got2 = m.find_matches(None, fr, (Rm, t), matcher)
    want2 = R.map_find_matches(pts, mp, has, exp >= 2, Rm, t, KITTI, fr)
    assert np.array_equal(got2, want2)
    assert len(m.find_matches(None, fr[:0], (Rm, t), matcher)) == 0          # empty frame (matcher.rs:29-31)
    behind = (np.eye(3), np.array([0.0, 0.0, -1e6]))
    assert len(m.find_matches(None, fr, behind, matcher)) == 0 and m.last_visible == 0


def test_full_size_map_store(orbx):
    """1 M map points x 5000 frame descriptors (BASELINE configs[3] through the reference's own call,
    map.rs:107): size-independent properties."""
    n, nf = 1_000_000, 5000
    rng = np.random.default_rng(9)
    mp, fr = synth.map_and_frame_descriptors(n, nf, f=2)
    pts = np.stack([rng.uniform(-50, 50, n), rng.uniform(-20, 20, n), rng.uniform(1, 100, n)], 1)
    m = orbx.Map(orbx.CameraIntrinsics(*KITTI), capacity=n)
    m.add_points(pts, mp)
    pose = (np.eye(3), np.zeros(3))
    matcher = orbx.FeatureMatcher()
    got = m.find_matches(None, fr, pose, matcher)
    _, rvis = R.map_project(pts, *pose, *KITTI)
    assert m.last_visible == int(rvis.sum())
    assert np.all(rvis[got[:, 0]]) and np.all(np.diff(got[:, 0]) > 0)
    # every pair is a true nearest neighbour under the filter threshold: re-derive the distances
    d = np.unpackbits(mp[got[:, 0]] ^ fr[got[:, 1]], axis=1).sum(1)
    assert d.max() < max(2.0 * d.min(), 30.0)
    # the noisy copies whose source row is visible must all be found (their distance is far below 30..60)
    src = {}
    pick = np.random.default_rng(8 + 2).choice(n, nf // 2, replace=False)    # generator of synth.map_and_frame_descriptors
    found = dict(zip(got[:, 0].tolist(), got[:, 1].tolist()))
    exact = [(int(p), k) for k, p in enumerate(pick) if rvis[p] and np.unpackbits(mp[p] ^ fr[k]).sum() < d.max()]
    assert all(found.get(p) == k for p, k in exact[:2000])
