"""GPU parity of the tensor-core Hamming kernel (csrc/match_mma.cu: tcgen05.mma kind::i8, accumulators in TMEM) against
the C oracle and against the XOR+POPC kernel: indices, distances, counts and the tie rule (lowest train index; lowest
query index in the reverse search) must be bit-identical on every shape - tiny, ragged, > 65536 rows, train sets that
are split across units, batched problems with per-problem device counts."""
import numpy as np
import pytest

from oracle import restated as R, synth, c_oracle

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def orbx():
    import slam_rs_b200 as s
    return s


def with_ties(nq, nt, seed):
    q, t = synth.random_descriptors(nq, seed), synth.random_descriptors(nt, seed + 100)
    if nt > 20 and nq > 4:
        t[nt // 2] = t[nt - 1] = t[3]        # duplicate train rows: forward ties, also across tiles
        q[0] = t[3]
        q[nq - 1] = q[2] = q[1]              # duplicate query rows: reverse ties
        t[5] = q[1]
        if nt > 600:
            t[255] = t[256] = t[511] = t[7]  # ties on both sides of a 256-row tile boundary
            q[3] = t[7]
    return q, t


@pytest.mark.parametrize("nq,nt,seed", [(1, 1, 21), (1, 2, 22), (7, 255, 23), (128, 256, 24), (129, 257, 25), (300, 1, 26),
                                        (2000, 2000, 27), (2336, 1987, 28), (5000, 5000, 29), (333, 70001, 30),
                                        (70001, 300, 31), (4999, 131073, 32)])
def test_mma_vs_oracle_and_popc(orbx, nq, nt, seed):
    q, t = with_ties(nq, nt, seed)
    mma, popc = orbx.FeatureMatcher(mode="mma"), orbx.FeatureMatcher(mode="popc")
    idx, dist, c = c_oracle.knn_mt(q, t, 2)
    knn, cnt = mma.knn_match(q, t)
    assert np.array_equal(cnt, c)
    for j in range(2):
        ok = cnt > j
        assert np.array_equal(knn["train_idx"][ok, j], idx[ok, j])
        assert np.array_equal(knn["distance"][ok, j], dist[ok, j].astype(np.float32))
        assert np.array_equal(knn["query_idx"][ok, j], np.nonzero(ok)[0])
    kp, cp = popc.knn_match(q, t)
    ok = cnt > 1
    assert np.array_equal(cnt, cp) and np.array_equal(knn[:, 0], kp[:, 0]) and np.array_equal(knn[ok], kp[ok])
    assert np.array_equal(mma.match_descriptors(q, t), popc.match_descriptors(q, t))


@pytest.mark.parametrize("nq,nt,seed", [(5, 300, 41), (130, 127, 42), (700, 900, 43), (300, 70001, 44), (70001, 300, 45),
                                        (2049, 2047, 46)])
def test_mma_cross_check(orbx, nq, nt, seed):
    """crossCheck: the reverse best-1 comes out of the same accumulators (CREDUX.MAX over the rows of a warp)."""
    q, t = with_ties(nq, nt, seed)
    fwd_i, fwd_d, c = c_oracle.knn_mt(q, t, 2)
    rev_i, _, _ = c_oracle.knn_mt(t, q, 1)
    mutual = rev_i[fwd_i[:, 0], 0] == np.arange(nq)
    mx = orbx.FeatureMatcher(cross_check=True, mode="mma")
    out = mx.match_descriptors(q, t)
    assert np.array_equal(out["query_idx"], np.nonzero(mutual)[0])
    assert np.array_equal(out["train_idx"], fwd_i[mutual, 0])
    assert np.array_equal(out["distance"], fwd_d[mutual, 0].astype(np.float32))
    knn, cnt = mx.knn_match(q, t)
    assert np.array_equal(cnt, np.where(mutual, c, 0))
    px = orbx.FeatureMatcher(cross_check=True, mode="popc")
    assert np.array_equal(out, px.match_descriptors(q, t))


def test_mma_batched_device_counts(orbx):
    """The frame-to-frame form (orbx_knn_match2_batch over consecutive sets of one array, per-problem device counts,
    cross-check, ratio test): the chained expansion of npairs + 1 sets, ragged counts incl. 0, 1 and cap."""
    import torch
    rng = np.random.default_rng(7)
    T, cap = 9, 640
    counts = np.array([640, 500, 0, 1, 257, 256, 129, 640, 3], np.int32)
    desc = rng.integers(0, 256, (T, cap, 32), dtype=np.uint8)
    for t in range(1, T):                       # real matches between consecutive sets
        k = min(counts[t], counts[t - 1]) // 2
        desc[t, :k] = desc[t - 1, :k]
        desc[t, :k, 5] ^= 1
    d = torch.from_numpy(desc).cuda()
    n = torch.from_numpy(counts).cuda()
    res = {}
    for mode in ("mma", "popc"):
        m = orbx.FeatureMatcher(cross_check=True, mode=mode)
        knn = torch.zeros((T, cap, 2, 16), dtype=torch.uint8, device="cuda")
        npq = torch.zeros((T, cap), dtype=torch.int32, device="cuda")
        good = torch.zeros((T, cap, 16), dtype=torch.uint8, device="cuda")
        ng = torch.zeros((T,), dtype=torch.int32, device="cuda")
        torch.cuda.synchronize()
        m.knn_match_batch_device(d, cap * 32, n, d[1:], cap * 32, n[1:], T - 1, cap, knn[1:], npq[1:])
        m.ratio_test_batch_device(knn[1:], npq[1:], n, T - 1, cap, 0.75, good[1:], ng[1:])
        m.sync()
        res[mode] = (knn.cpu().numpy(), npq.cpu().numpy(), good.cpu().numpy(), ng.cpu().numpy())
    for t in range(1, T):
        nq, nt = counts[t - 1], counts[t]
        kn = res["mma"][0][t].view(orbx.DMATCH_DTYPE).reshape(cap, 2)[:nq]
        cn = res["mma"][1][t][:nq]
        idx, dist, c = R.knn(desc[t - 1, :nq], desc[t, :nt], 2)
        rev, _, _ = R.knn(desc[t, :nt], desc[t - 1, :nq], 1)
        mutual = (rev[idx[:, 0], 0] == np.arange(nq)) if nt else np.zeros(nq, bool)
        assert np.array_equal(cn, np.where(mutual, c, 0)), t
        ok = cn > 0
        assert np.array_equal(kn["train_idx"][ok, 0], idx[ok, 0]) and np.array_equal(kn["distance"][ok, 0], dist[ok, 0])
        ok2 = cn > 1
        assert np.array_equal(kn["train_idx"][ok2, 1], idx[ok2, 1])
        # the two kernels agree on everything the API defines (counts, records of counted rows, good lists)
        assert np.array_equal(res["mma"][1][t][:nq], res["popc"][1][t][:nq])
        assert res["mma"][3][t] == res["popc"][3][t]
        g = res["mma"][3][t]
        assert np.array_equal(res["mma"][2][t][:g], res["popc"][2][t][:g])


def test_mma_full_size_map_all_queries(orbx):
    """BASELINE configs[3] at full size - 5000 frame descriptors against 1 M map descriptors - checked on ALL 5000
    queries against the C oracle (the train set is cut into 15 unit ranges whose partial keys are merged)."""
    mp, fr = synth.map_and_frame_descriptors(1_000_000, 5000, f=1)
    idx, dist, c = c_oracle.knn_mt(fr, mp, 2)
    for mode in ("mma", "popc"):
        knn, cnt = orbx.FeatureMatcher(mode=mode).knn_match(fr, mp)
        assert np.array_equal(cnt, c), mode
        assert np.array_equal(knn["train_idx"], idx) and np.array_equal(knn["distance"], dist.astype(np.float32)), mode
        assert np.array_equal(knn["query_idx"][:, 0], np.arange(5000))
