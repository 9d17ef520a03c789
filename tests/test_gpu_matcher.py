"""GPU parity: FeatureMatcher through the C ABI vs the oracle (bit-exact indices and distances)."""
import os

import numpy as np
import pytest

from oracle import restated as R, synth, cv_ref, c_oracle
from tests.test_oracle_cpu import GOLD
from tests.util import compare_orb

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def orbx():
    import slam_rs_b200 as s
    return s


def check_knn(out, cnt, idx, dist, c, q0=0):
    assert np.array_equal(cnt, c)
    for j in range(out.shape[1]):
        ok = cnt > j
        assert np.array_equal(out["train_idx"][ok, j], idx[ok, j])
        assert np.array_equal(out["distance"][ok, j], dist[ok, j].astype(np.float32))
        assert np.array_equal(out["query_idx"][ok, j], np.nonzero(ok)[0] + q0)
        assert np.all(out["img_idx"][ok, j] == 0)


def test_golden_match_vectors(orbx):
    g = np.load(os.path.join(GOLD, "match_700x900.npz"))
    q, t = g["q"], g["t"]
    m = orbx.FeatureMatcher()
    out = m.match_descriptors(q, t)
    assert np.array_equal(out["query_idx"], g["match"][:, 0].astype(np.int32))
    assert np.array_equal(out["train_idx"], g["match"][:, 1].astype(np.int32))
    assert np.array_equal(out["distance"], g["match"][:, 3].astype(np.float32))
    knn, cnt = m.knn_match(q, t)
    assert np.all(cnt == 2)
    assert np.array_equal(knn["train_idx"], g["knn2"][:, :, 1].astype(np.int32))
    assert np.array_equal(knn["distance"], g["knn2"][:, :, 3].astype(np.float32))
    mx = orbx.FeatureMatcher(cross_check=True).match_descriptors(q, t)
    assert np.array_equal(mx["query_idx"], g["cross"][:, 0].astype(np.int32))
    assert np.array_equal(mx["train_idx"], g["cross"][:, 1].astype(np.int32))
    assert np.array_equal(mx["distance"], g["cross"][:, 3].astype(np.float32))


def test_empty_and_tiny(orbx):
    # matcher.rs:79-86: empty x empty -> Ok, 0 matches; matcher.rs:29-31: either side empty -> empty
    m = orbx.FeatureMatcher()
    e = np.zeros((0, 32), np.uint8)
    q = synth.random_descriptors(5, 1)
    assert len(m.match_descriptors(e, e)) == 0
    assert len(m.match_descriptors(q, e)) == 0
    assert len(m.match_descriptors(e, q)) == 0
    knn, cnt = m.knn_match(q, e)
    assert np.all(cnt == 0)
    # one train row: k=2 yields a single neighbour per query
    t1 = synth.random_descriptors(1, 2)
    knn, cnt = m.knn_match(q, t1)
    assert np.all(cnt == 1) and np.all(knn["train_idx"][:, 0] == 0)
    assert np.array_equal(knn["distance"][:, 0], R.hamming_matrix(q, t1)[:, 0].astype(np.float32))
    # identical rows everywhere: every tie resolves to the lowest train index
    t = np.repeat(synth.random_descriptors(1, 3), 300, axis=0)
    knn, cnt = m.knn_match(q, t)
    assert np.all(knn["train_idx"][:, 0] == 0) and np.all(knn["train_idx"][:, 1] == 1)
    with pytest.raises(ValueError):
        m.match_descriptors(np.zeros((3, 16), np.uint8), t)


@pytest.mark.parametrize("nq,nt,seed", [(1, 1, 1), (127, 255, 2), (129, 257, 3), (2000, 2000, 4), (5000, 5000, 5),
                                        (333, 70001, 6)])
def test_knn_vs_oracle_sizes(orbx, nq, nt, seed):
    q, t = synth.random_descriptors(nq, seed), synth.random_descriptors(nt, seed + 100)
    if nt > 20:
        t[nt // 2] = t[nt - 1] = t[3]
        q[0] = t[3]
    m = orbx.FeatureMatcher()
    idx, dist, c = c_oracle.knn(q, t, 2)
    knn, cnt = m.knn_match(q, t)
    check_knn(knn, cnt, idx, dist, c)
    out = m.match_descriptors(q, t)
    assert np.array_equal(out["train_idx"], idx[:, 0]) and np.array_equal(out["distance"], dist[:, 0].astype(np.float32))


def test_orb_descriptors_match_and_filters(orbx):
    seq = synth.sequence(376, 1241, 2)
    det = orbx.OrbDetector(2000, max_width=1241, max_height=376)
    (k0, d0), (k1, d1) = det.detect_and_compute(seq[0]), det.detect_and_compute(seq[1])
    m = orbx.FeatureMatcher()
    # reference loop: match (k=1) + filter_good_matches(2.0)  (visual_odometry.rs:106-107)
    raw = m.match_descriptors(d0, d1)
    ref = R.match_k1(d0, d1)
    assert np.array_equal(raw["train_idx"], ref[:, 1]) and np.array_equal(raw["distance"], ref[:, 2].astype(np.float32))
    good = m.filter_good_matches(raw, 2.0)
    rgood = R.filter_good_matches(ref, 2.0)
    assert np.array_equal(good["query_idx"], rgood[:, 0]) and np.array_equal(good["train_idx"], rgood[:, 1])
    assert len(good) > 300
    # the true motion is a 3 px shift: filtered matches must agree with it
    dx = k0["x"][good["query_idx"]] - k1["x"][good["train_idx"]]
    assert np.mean(np.abs(dx - 3.0) < 2.5) > 0.9
    # north_star composition: k=2 + Lowe ratio + cross-check
    knn, cnt = m.knn_match(d0, d1)
    idx, dist, c = R.knn(d0, d1, 2)
    check_knn(knn, cnt, idx, dist, c)
    kept = m.ratio_test(knn, cnt, 0.75)
    rk = R.ratio_test(idx, dist, c, 0.75)
    assert np.array_equal(kept["query_idx"], np.nonzero(rk)[0]) and np.array_equal(kept["train_idx"], idx[rk, 0])
    mx = orbx.FeatureMatcher(cross_check=True)
    knn_x, cnt_x = mx.knn_match(d0, d1)
    mutual = np.zeros(len(d0), bool)
    mutual[R.cross_check(d0, d1)[:, 0]] = True
    assert np.array_equal(cnt_x, np.where(mutual, c, 0))
    kept_x = mx.ratio_test(knn_x, cnt_x, 0.75)
    assert np.array_equal(kept_x["query_idx"], np.nonzero(rk & mutual)[0])
    if cv_ref.HAVE_CV2:
        assert np.array_equal(cv_ref.match_k1(d0, d1), ref)
        x = cv_ref.match_k1(d0, d1, cross_check=True)
        assert np.array_equal(mx.match_descriptors(d0, d1)["query_idx"], x[:, 0])


def test_map_to_frame_both_orientations(orbx):
    """Config C4 shape, scaled to what the C oracle checks in seconds, in both orientations
    (reference: query = map, train = frame — map.rs:107; north_star: query = frame, train = map)."""
    mp, fr = synth.map_and_frame_descriptors(120000, 1000, f=0)
    m = orbx.FeatureMatcher()
    out = m.match_descriptors(mp, fr)
    idx, dist, _ = c_oracle.knn(mp, fr, 1)
    assert np.array_equal(out["train_idx"], idx[:, 0]) and np.array_equal(out["distance"], dist[:, 0].astype(np.float32))
    knn, cnt = m.knn_match(fr, mp)              # > 65536 train rows: split + merge path
    idx, dist, c = c_oracle.knn(fr, mp, 2)
    check_knn(knn, cnt, idx, dist, c)


def test_full_size_map_properties(orbx):
    """1M map x 5000 frame descriptors (BASELINE config 3 per-frame problem): size-independent properties +
    an oracle check on a slice of the queries."""
    mp, fr = synth.map_and_frame_descriptors(1_000_000, 5000, f=1)
    m = orbx.FeatureMatcher()
    knn, cnt = m.knn_match(fr, mp)
    assert np.all(cnt == 2)
    assert np.all(knn["distance"][:, 0] <= knn["distance"][:, 1])
    tie = knn["distance"][:, 0] == knn["distance"][:, 1]
    assert np.all(knn["train_idx"][tie, 0] < knn["train_idx"][tie, 1])
    # noisy copies (first half) must find their source row: distance ~ Binomial(256, 0.1)
    d = np.unpackbits(fr[:2500] ^ mp[knn["train_idx"][:2500, 0]], axis=1).sum(1)
    assert np.array_equal(d.astype(np.float32), knn["distance"][:2500, 0])
    assert d.max() < 60 and knn["distance"][2500:, 0].min() > 60
    sl = slice(2400, 2600)
    idx, dist, c = c_oracle.knn(fr[sl], mp, 2)
    check_knn(knn[sl], cnt[sl], idx, dist, c, q0=sl.start)
    # linearity / symmetry: matching the best rows back gives distance 0 to themselves
    back = m.match_descriptors(mp[knn["train_idx"][:64, 0]], mp)
    assert np.all(back["distance"] == 0)


def test_sequence_front_end_matches_per_frame_api(orbx):
    import torch
    from slam_rs_b200.frontend import SequenceFrontEnd
    T, h, w, nf = 7, 200, 320, 400
    seq = synth.sequence(h, w, T, canvas_seed=77)
    fe = SequenceFrontEnd(w, h, nfeatures=nf, max_frames=T, chunk=3, reference_filter=True)
    dev = torch.zeros((T, h, fe.pitch), dtype=torch.uint8, device="cuda")
    dev[:, :, :w] = torch.from_numpy(seq).cuda()
    torch.cuda.synchronize()
    fe.run_device(dev)
    torch.cuda.synchronize()
    det = orbx.OrbDetector(nf, max_width=w, max_height=h)
    per = [det.detect_and_compute(f) for f in seq]
    for t in range(T):
        k, d = fe.frame_result(t)
        assert np.array_equal(k, per[t][0]) and np.array_equal(d, per[t][1])
    for t in range(1, T):
        d0, d1 = per[t - 1][1], per[t][1]
        idx, dist, c = R.knn(d0, d1, 2)
        mutual = np.zeros(len(d0), bool)
        mutual[R.cross_check(d0, d1)[:, 0]] = True
        keep = R.ratio_test(idx, dist, c, 0.75) & mutual
        good = fe.pair_result(t)
        assert np.array_equal(good["query_idx"], np.nonzero(keep)[0])
        assert np.array_equal(good["train_idx"], idx[keep, 0])
        assert np.array_equal(good["distance"], dist[keep, 0].astype(np.float32))
        n = int(fe.n_ref_good[t].item())
        rg = fe.ref_good[t, :n].cpu().numpy().view(orbx.DMATCH_DTYPE).reshape(-1)
        ref = R.filter_good_matches(R.match_k1(d0, d1), 2.0)
        assert np.array_equal(rg["query_idx"], ref[:, 0]) and np.array_equal(rg["train_idx"], ref[:, 1])
    # the host (end-to-end) path returns the same bytes
    out = fe.alloc_host_outputs()
    fe.run_host(torch.from_numpy(seq).pin_memory(), out)
    torch.cuda.synchronize()
    for t in range(T):
        n = int(out["n"][t])
        assert np.array_equal(out["desc"][t, :n].numpy(), per[t][1])
    assert np.array_equal(out["n_good"][1:T].numpy(), fe.n_good[1:T].cpu().numpy())
    # ... and so does the two-stream pipeline (stages alternate between two detector / matcher handles)
    fe2 = SequenceFrontEnd(w, h, nfeatures=nf, max_frames=T, chunk=T, host_chunk=2, dual=True)
    out2 = fe2.alloc_host_outputs()
    fe2.run_host(torch.from_numpy(seq).pin_memory(), out2)
    torch.cuda.synchronize()
    for key in ("n", "n_good"):
        assert np.array_equal(out2[key].numpy(), out[key].numpy())
    for t in range(T):
        n, g = int(out["n"][t]), int(out["n_good"][t])
        assert np.array_equal(out2["desc"][t, :n].numpy(), out["desc"][t, :n].numpy())
        assert np.array_equal(out2["kps"][t, :n].numpy(), out["kps"][t, :n].numpy())
        if t:
            assert np.array_equal(out2["good"][t, :g].numpy(), out["good"][t, :g].numpy())


def test_map_to_frame_matcher_class(orbx):
    """slam_rs_b200.mapmatch.MapToFrameMatcher at world == 1 (the multi-rank slicing/gather logic is covered
    on CPU by tests/test_abi_cpu.py::test_query_sharding_world2_gloo): both orientations vs the C oracle,
    plus a hand-sliced query range to check the global query_idx offset."""
    import torch
    from slam_rs_b200.mapmatch import MapToFrameMatcher
    from slam_rs_b200.sharding import query_slice
    mp, fr = synth.map_and_frame_descriptors(70000, 601, f=2)
    mm = MapToFrameMatcher(device=0)
    assert mm.set_train(mp) == len(mp)
    rec, cnt = mm.knn2(torch.from_numpy(fr).cuda())
    torch.cuda.synchronize()
    knn = rec.cpu().numpy().view(orbx.DMATCH_DTYPE).reshape(len(fr), 2)
    idx, dist, c = c_oracle.knn(fr, mp, 2)
    check_knn(knn, cnt.cpu().numpy(), idx, dist, c)
    # a rank-1-of-2 style slice: records must carry global query indices
    mm.rank, mm.world = 1, 2
    lo, hi = query_slice(len(fr), 1, 2)
    mm_gather = mm._gather
    mm._gather = lambda part, total: part          # no process group here: look at the local part only
    rec1, cnt1 = mm.knn2(torch.from_numpy(fr).cuda(), replicate_query=False)
    torch.cuda.synchronize()
    part = rec1.cpu().numpy().view(orbx.DMATCH_DTYPE).reshape(hi - lo, 2)
    check_knn(part, cnt1.cpu().numpy(), idx[lo:hi], dist[lo:hi], c[lo:hi], q0=lo)
    mm.rank, mm.world, mm._gather = 0, 1, mm_gather
    # reference orientation (map.rs:107-108): query = map, train = frame, k = 1 + filter_good_matches(2.0)
    good = mm.match_train_as_query(torch.from_numpy(fr).cuda(), 2.0)
    ref = R.filter_good_matches(R.match_k1(mp, fr), 2.0) if len(mp) <= 5000 else None
    i1, d1, _ = c_oracle.knn(mp, fr, 1)
    thr = max(2.0 * float(d1.min()), 30.0)
    keep = np.nonzero(d1[:, 0] < thr)[0]
    assert np.array_equal(good["query_idx"], keep) and np.array_equal(good["train_idx"], i1[keep, 0])
    assert np.array_equal(good["distance"], d1[keep, 0].astype(np.float32))


@pytest.mark.parametrize("nq,nt,seed", [(1, 1, 11), (5, 300, 12), (130, 127, 13), (700, 900, 14), (300, 70001, 15),
                                        (70001, 300, 16), (2049, 2047, 17)])
def test_cross_check_fused_vs_oracle(orbx, nq, nt, seed):
    """crossCheck=true (mutual first-arg-min, ascending queryIdx — SURVEY.md Appendix A S10): the reverse search is
    produced by the same kernel launch as the forward one (CREDUX column minima), including the train-split path
    (> 65536 train rows) and more than 65536 queries; forced duplicate rows exercise the lowest-index tie rule in
    both directions."""
    q, t = synth.random_descriptors(nq, seed), synth.random_descriptors(nt, seed + 100)
    if nt > 20 and nq > 4:
        t[nt // 2] = t[nt - 1] = t[3]        # duplicate train rows: forward ties
        q[0] = t[3]
        q[nq - 1] = q[2] = q[1]              # duplicate query rows: reverse ties
        t[5] = q[1]
    mx = orbx.FeatureMatcher(cross_check=True)
    fwd_i, fwd_d, c = c_oracle.knn(q, t, 2)
    rev_i, _, _ = c_oracle.knn(t, q, 1)
    mutual = rev_i[fwd_i[:, 0], 0] == np.arange(nq)
    out = mx.match_descriptors(q, t)
    assert np.array_equal(out["query_idx"], np.nonzero(mutual)[0])
    assert np.array_equal(out["train_idx"], fwd_i[mutual, 0])
    assert np.array_equal(out["distance"], fwd_d[mutual, 0].astype(np.float32))
    knn, cnt = mx.knn_match(q, t)
    assert np.array_equal(cnt, np.where(mutual, c, 0))
    ok = cnt > 0
    assert np.array_equal(knn["train_idx"][ok, 0], fwd_i[ok, 0])
    if cv_ref.HAVE_CV2 and nt < 262144 and nq < 5000:
        x = cv_ref.match_k1(q, t, cross_check=True)
        assert np.array_equal(out["query_idx"], x[:, 0]) and np.array_equal(out["train_idx"], x[:, 1])


def test_extract_matched_points(orbx):
    """PoseEstimator::extract_matched_points (pose.rs:27-44) through the C ABI: host arrays, the batched device
    form inside the sequence front end, the empty case and the reference's out-of-range error."""
    import torch
    from slam_rs_b200.frontend import SequenceFrontEnd
    T, h, w, nf = 5, 200, 320, 400
    seq = synth.sequence(h, w, T, canvas_seed=91)
    fe = SequenceFrontEnd(w, h, nfeatures=nf, max_frames=T, chunk=T, points=True)
    dev = torch.zeros((T, h, fe.pitch), dtype=torch.uint8, device="cuda")
    dev[:, :, :w] = torch.from_numpy(seq).cuda()
    fe.run_device(dev)
    torch.cuda.synchronize()
    assert int(fe.bad_index.item()) == 0
    pe = orbx.PoseEstimator(orbx.CameraIntrinsics(718.856, 718.856, 607.1928, 185.2157))
    assert pe.min_matches == 8
    total = 0
    for t in range(1, T):
        k0, _ = fe.frame_result(t - 1)
        k1, _ = fe.frame_result(t)
        good = fe.pair_result(t)
        xy0 = np.stack([k0["x"], k0["y"]], 1)
        xy1 = np.stack([k1["x"], k1["y"]], 1)
        want1, want2 = R.extract_matched_points(xy0, xy1, np.stack([good["query_idx"], good["train_idx"]], 1))
        got1, got2 = fe.pair_points(t)                       # batched device form
        assert np.array_equal(got1, want1) and np.array_equal(got2, want2)
        h1, h2 = pe.extract_matched_points(k0, k1, good)     # host-array form, reference method name
        assert np.array_equal(h1, want1) and np.array_equal(h2, want2)
        total += len(good)
    assert total > 50
    # empty match list -> two empty lists
    e1, e2 = pe.extract_matched_points(k0, k1, good[:0])
    assert e1.shape == (0, 2) and e2.shape == (0, 2)
    # an index past the keypoint array is an error in the reference (`kp1.get(idx)?`)
    bad = good[:3].copy()
    bad["train_idx"][1] = len(k1)
    with pytest.raises(orbx.OrbxError):
        pe.extract_matched_points(k0, k1, bad)
    bad["train_idx"][1] = -1
    with pytest.raises(orbx.OrbxError):
        pe.extract_matched_points(k0, k1, bad)


@pytest.mark.skipif(not cv_ref.HAVE_CV2, reason="cv2 not importable")
def test_kitti_shaped_sequence_full_size(orbx):
    """BASELINE configs[1] at full frame size (1241x376, nfeatures 2000) on a 48-frame stretch of the bench's own
    sequence, through every front-end path (one launch set, chunked, two-stream host pipeline): sampled frames
    and pairs are compared with cv2 (the OpenCV code the reference binds), all frames / pairs with the
    size-independent properties of the path."""
    import torch
    from slam_rs_b200.frontend import SequenceFrontEnd
    T, h, w, nf = 48, 376, 1241, 2000
    seq = synth.sequence(h, w, T)
    fe = SequenceFrontEnd(w, h, nfeatures=nf, max_frames=T, chunk=T, host_chunk=10, dual=True, points=True)
    dev = torch.zeros((T, h, fe.pitch), dtype=torch.uint8, device="cuda")
    dev[:, :, :w] = torch.from_numpy(seq).cuda()
    fe.run_device(dev)
    torch.cuda.synchronize()
    n = fe.n.cpu().numpy()
    n_good = fe.n_good.cpu().numpy()
    assert np.all(n == nf) and int(fe.bad_index.item()) == 0       # dense texture: every quota is filled
    desc_all = fe.desc.cpu().numpy()
    for t in (0, 17, T - 1):                                         # extraction vs cv2, bit for bit
        k, d = fe.frame_result(t)
        st = compare_orb(k, d, cv_ref.orb_detect_and_compute(seq[t], nf), h, w)
        assert st["missing"] == 0 and st["extra"] == 0 and st["desc_bits_diff"] == 0, st
        assert st["resp_eq"] == st["angle_eq"] == st["pt_eq"] == st["common"] == nf, st
    for t in (1, 30):                                                # matching vs cv2 BFMatcher on the same rows
        d0, d1 = desc_all[t - 1, :nf], desc_all[t, :nf]
        idx, dist, c = cv_ref.knn2(d0, d1)
        mutual = np.zeros(nf, bool)
        mutual[cv_ref.match_k1(d0, d1, cross_check=True)[:, 0]] = True
        keep = R.ratio_test(idx, dist, c, 0.75) & mutual
        good = fe.pair_result(t)
        assert np.array_equal(good["query_idx"], np.nonzero(keep)[0])
        assert np.array_equal(good["train_idx"], idx[keep, 0])
        assert np.array_equal(good["distance"], dist[keep, 0].astype(np.float32))
    for t in range(1, T):                                            # properties of every pair
        g = fe.pair_result(t)
        assert len(g) == n_good[t] > 0
        assert np.all(np.diff(g["query_idx"]) > 0)                   # query order, one match per query
        assert len(np.unique(g["train_idx"])) == len(g)              # cross-check => injective
        q, tr = desc_all[t - 1][g["query_idx"]], desc_all[t][g["train_idx"]]
        assert np.array_equal(np.unpackbits(q ^ tr, axis=1).sum(1), g["distance"].astype(np.int64))
    # chunked launch sets and the two-stream host pipeline return the same bytes
    ref = {k: getattr(fe, k).cpu().numpy().copy() for k in ("n", "kps", "desc", "good", "n_good")}
    fe2 = SequenceFrontEnd(w, h, nfeatures=nf, max_frames=T, chunk=7)
    fe2.run_device(dev)
    out = fe.alloc_host_outputs()
    fe.run_host(torch.from_numpy(seq).pin_memory(), out)
    torch.cuda.synchronize()
    for t in range(T):
        for name, m in (("kps", n[t]), ("desc", n[t]), ("good", n_good[t] if t else 0)):
            a = ref[name][t, :m]
            assert np.array_equal(getattr(fe2, name)[t, :m].cpu().numpy(), a), (name, t)
            assert np.array_equal(out[name][t, :m].numpy(), a), (name, t)
    assert np.array_equal(out["n"].numpy(), n) and np.array_equal(out["n_good"][1:].numpy(), n_good[1:])


def test_hd_pipeline_fresh_handles_host_first(orbx):
    """1920x1080, nfeatures=5000 (BASELINE configs[2] shape) through the two-stream host pipeline on FRESH handles
    while the copy engine is busy with the queued frame uploads: the first launch of a new detector must already see
    its resize / FAST tables (regression: the tables used to be uploaded by a legacy-stream cudaMemcpy from pageable
    memory, which is not ordered against the non-blocking stream the kernels run on).  Level 0 has pitch == width
    here (1920 is a multiple of 128), unlike the 1241-wide case."""
    import torch
    from slam_rs_b200.frontend import SequenceFrontEnd
    T, h, w, nf = 48, 1080, 1920, 5000
    seq = synth.sequence(h, w, T, canvas_seed=201)
    pinned = torch.from_numpy(seq).pin_memory()
    fe = SequenceFrontEnd(w, h, nfeatures=nf, max_frames=T, chunk=T, host_chunk=12, dual=True)
    out = fe.alloc_host_outputs()
    fe.run_host(pinned, out)                                                   # first use of both detector handles
    torch.cuda.synchronize()
    fe2 = SequenceFrontEnd(w, h, nfeatures=nf, max_frames=T, chunk=T)
    dev = torch.zeros((T, h, fe2.pitch), dtype=torch.uint8, device="cuda")
    dev[:, :, :w].copy_(pinned)
    fe2.run_device(dev)
    torch.cuda.synchronize()
    assert np.array_equal(out["n"].numpy(), fe2.n.cpu().numpy())
    assert np.array_equal(out["n_good"][1:].numpy(), fe2.n_good[1:].cpu().numpy())
    for t in range(T):
        n, g = int(out["n"][t]), int(out["n_good"][t])
        assert np.array_equal(out["kps"][t, :n].numpy(), fe2.kps[t, :n].cpu().numpy())
        assert np.array_equal(out["desc"][t, :n].numpy(), fe2.desc[t, :n].cpu().numpy())
        if t:
            assert np.array_equal(out["good"][t, :g].numpy(), fe2.good[t, :g].cpu().numpy())
    for t in (0, 3, 13, T - 1):                                                # first frame of a handle's first batch too
        n = int(out["n"][t])
        k = out["kps"][t, :n].numpy().view(orbx.KEYPOINT_DTYPE).reshape(-1)
        ref = cv_ref.orb_detect_and_compute(seq[t], nf)
        st = compare_orb(k, out["desc"][t, :n].numpy(), ref, h, w)
        assert st["missing"] == 0 and st["extra"] == 0 and st["desc_bits_diff"] == 0, st
