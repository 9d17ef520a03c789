"""GPU: the native sequence front end (csrc/sequence.cu, orbx_sequence_*) against the Python-scheduled one
(slam-rs_b200/frontend.py::SequenceFrontEnd, itself checked against the per-frame API and the oracle in
tests/test_gpu_matcher.py): same kernels, same launch sets, so every output must be identical bit for bit.

"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _valid_rows_equal(a, b, counts, what):
    for t, c in enumerate(counts):
        assert np.array_equal(a[t, :c], b[t, :c]), (what, t)


@pytest.mark.parametrize("host_chunk", [3, 16])
def test_native_equals_python_front_end(host_chunk):
    import torch
    from oracle import synth
    from slam_rs_b200.frontend import SequenceFrontEnd, NativeSequence
    T, h, w, nf = 11, 376, 1241, 600
    seq = synth.sequence(h, w, T, canvas_seed=51)
    fe = SequenceFrontEnd(w, h, nfeatures=nf, max_frames=T, chunk=T, host_chunk=host_chunk, dual=True)
    want = fe.alloc_host_outputs()
    pinned = torch.from_numpy(seq).pin_memory()
    fe.run_host(pinned, want)
    torch.cuda.synchronize()

    ns = NativeSequence(w, h, nfeatures=nf, max_frames=T, chunk=T, host_chunk=host_chunk)
    assert ns.cap == fe.cap
    got = ns.alloc_host_outputs()
    for _ in range(2):                                                         # the second pass reuses streams / events
        ns.run_host(pinned, got)
        ns.sync()
        n = want["n"].numpy()
        assert np.array_equal(got["n"].numpy(), n) and n.min() > 0
        _valid_rows_equal(got["kps"].numpy(), want["kps"].numpy(), n, "kps")
        _valid_rows_equal(got["desc"].numpy(), want["desc"].numpy(), n, "desc")
        ng = want["n_good"].numpy()
        assert np.array_equal(got["n_good"].numpy()[1:], ng[1:]) and ng[1:].min() > 0
        _valid_rows_equal(got["good"].numpy()[1:], want["good"].numpy()[1:], ng[1:], "good")

    # device-resident form: pitched frames in, results read back from the library's device buffers
    dev = torch.zeros((T, h, fe.pitch), dtype=torch.uint8, device="cuda")
    dev[:, :, :w] = torch.from_numpy(seq).cuda()
    torch.cuda.synchronize()
    ns.run_device(dev)
    res = ns.results_to_host(T)
    assert np.array_equal(res["n"], n)
    _valid_rows_equal(res["desc"], want["desc"].numpy(), n, "desc (device pass)")
    assert np.array_equal(res["n_good"][1:], ng[1:])
    _valid_rows_equal(res["good"][1:], want["good"].numpy()[1:], ng[1:], "good (device pass)")


def test_native_pipelined_passes_are_identical_to_serial_ones():
    """Consecutive run_host calls with the same T are pipelined into each other (chunk-level events instead of a join
    between the passes): every pass must return exactly what it returns when the passes are separated by a sync, also
    with different sequences in flight, a device pass or another T in between, and wait_pass(1) must cover the call
    before the most recent one."""
    import torch
    from oracle import synth
    import slam_rs_b200 as s
    from slam_rs_b200.frontend import NativeSequence
    T, h, w, nf = 26, 240, 320, 400
    seqs = [torch.from_numpy(synth.sequence(h, w, T, canvas_seed=70 + k, noise_seed0=500 * k)).pin_memory() for k in range(3)]
    ns = NativeSequence(w, h, nfeatures=nf, max_frames=T, chunk=T, host_chunk=4)   # 8 chunks per pass
    keys = ("n", "kps", "desc", "n_good", "good")

    def snapshot(out):
        return {k: out[k].numpy().copy() for k in keys}

    def same(got, want, what):
        n, ng = want["n"], want["n_good"]
        assert np.array_equal(got["n"], n), what
        _valid_rows_equal(got["kps"], want["kps"], n, what + " kps")
        _valid_rows_equal(got["desc"], want["desc"], n, what + " desc")
        assert np.array_equal(got["n_good"][1:], ng[1:]), what
        _valid_rows_equal(got["good"][1:], want["good"][1:], ng[1:], what + " good")

    serial = []
    for k in range(3):                                                         # serial: one pass, one sync
        out = ns.alloc_host_outputs()
        ns.run_host(seqs[k], out)
        ns.sync()
        serial.append(snapshot(out))
        dev = seqs[k].cuda()
        ns.run_device(dev)                                                     # a device pass ends the pipelining (join)
        ns.sync()
    assert not np.array_equal(serial[0]["desc"], serial[1]["desc"])

    outs = [ns.alloc_host_outputs() for _ in range(3)]
    order = [0, 1, 2, 1, 0, 2, 2, 0]
    for rep in range(3):
        for i, k in enumerate(order):                                          # back to back, no sync in between
            ns.run_host(seqs[k], outs[i % 3])
            if i >= 1:
                ns.wait_pass(1)                                                # the call before this one has finished
                same(snapshot(outs[(i - 1) % 3]), serial[order[i - 1]], "pipelined pass %d of round %d" % (i - 1, rep))
        ns.wait_pass(0)
        same(snapshot(outs[(len(order) - 1) % 3]), serial[order[-1]], "last pass of round %d" % rep)
        if rep == 0:                                                           # a shorter sequence in between: joins
            short = ns.alloc_host_outputs()
            ns.run_host(seqs[0][:9], short)
            ns.sync()
            got = snapshot(short)
            assert np.array_equal(got["n"][:9], serial[0]["n"][:9])
            _valid_rows_equal(got["desc"][:9], serial[0]["desc"][:9], serial[0]["n"][:9], "short pass")
    ns.sync()
    # different lengths back to back without a sync: the chunk-level waits go by frame range, not by chunk index
    short = ns.alloc_host_outputs()
    ns.run_host(seqs[0], outs[0])
    ns.run_host(seqs[1][:9], short)
    ns.run_host(seqs[2], outs[2])
    ns.run_host(seqs[0][:17], outs[1])
    ns.sync()
    same(snapshot(outs[0]), serial[0], "long pass before a short one")
    same(snapshot(outs[2]), serial[2], "long pass after a short one")
    for got, k, n9 in ((snapshot(short), 1, 9), (snapshot(outs[1]), 0, 17)):
        assert np.array_equal(got["n"][:n9], serial[k]["n"][:n9])
        _valid_rows_equal(got["desc"][:n9], serial[k]["desc"][:n9], serial[k]["n"][:n9], "short pipelined pass desc")
        assert np.array_equal(got["n_good"][1:n9], serial[k]["n_good"][1:n9])
        _valid_rows_equal(got["good"][1:n9], serial[k]["good"][1:n9], serial[k]["n_good"][1:n9], "short pipelined pass good")
    with pytest.raises(s.OrbxError):
        ns.wait_pass(5)
    ns.wait_pass(2)
    ns.close()


def test_native_bad_arguments():
    import slam_rs_b200 as s
    from slam_rs_b200.frontend import NativeSequence
    with pytest.raises(s.OrbxError):
        NativeSequence(8, 8, max_frames=4)                                     # below the 16 px minimum
    ns = NativeSequence(320, 240, nfeatures=100, max_frames=4, chunk=4, host_chunk=2)
    import torch
    with pytest.raises(s.OrbxError):
        ns.run_device(torch.zeros((5, 240, 320), dtype=torch.uint8, device="cuda"))   # T > max_frames


def test_native_reference_filter_mode_and_timing():
    """orbx_sequence_set_reference_filter: the pairs are matched exactly as the reference's loop does
    (examples/visual_odometry.rs:106-107: match_descriptors k=1, no cross-check, then filter_good_matches(2.0)),
    pinned to the oracle's restatement of matcher.rs:24-66; orbx_sequence_elapsed reports the device time of the passes."""
    import torch
    from oracle import synth, restated as R
    import slam_rs_b200 as orbx
    from slam_rs_b200.frontend import NativeSequence
    T, h, w, nf = 6, 240, 320, 300
    seq = synth.sequence(h, w, T, canvas_seed=61)
    ns = NativeSequence(w, h, nfeatures=nf, max_frames=T, chunk=T, host_chunk=2)
    ns.set_reference_filter(True, 2.0)
    out = ns.alloc_host_outputs()
    ns.set_timing(True)
    pinned = torch.from_numpy(seq).pin_memory()
    for _ in range(3):
        ns.run_host(pinned, out)
    ns.sync()
    ms, passes = ns.elapsed()
    assert passes == 3 and 0 < ms < 10_000
    n, ng = out["n"].numpy(), out["n_good"].numpy()
    desc = out["desc"].numpy()
    good = out["good"].numpy().view(orbx.DMATCH_DTYPE).reshape(T, ns.cap)
    for t in range(1, T):
        want = R.filter_good_matches(R.match_k1(desc[t - 1, :n[t - 1]], desc[t, :n[t]]), 2.0)
        got = good[t, :ng[t]]
        assert ng[t] == len(want) > 0
        assert np.array_equal(got["query_idx"], want[:, 0]) and np.array_equal(got["train_idx"], want[:, 1])
        assert np.array_equal(got["distance"], want[:, 2].astype(np.float32))
    # back to the north_star composition on the same handle
    ns.set_reference_filter(False, 2.0)
    ns.run_host(pinned, out)
    ns.sync()
    m = orbx.FeatureMatcher(cross_check=True)
    knn, cnt = m.knn_match(desc[0, :n[0]], desc[1, :n[1]])
    lowe = m.ratio_test(knn, cnt, 0.75)
    g1 = out["good"].numpy().view(orbx.DMATCH_DTYPE).reshape(T, ns.cap)[1, :out["n_good"].numpy()[1]]
    assert np.array_equal(g1, lowe)


def test_native_sequence_reports_tie_flood_truncation():
    """A frame with more keypoints than the sequence's row capacity (a tie flood, see test_gpu_detector.py) is truncated
    to the capacity - n never exceeds it - and orbx_sequence_sync returns ORBX_E_CAPACITY once."""
    import torch
    from oracle import synth
    import slam_rs_b200 as orbx
    from slam_rs_b200.frontend import NativeSequence
    from tests.test_gpu_detector import dot_lattice
    T, h, w, nf = 4, 240, 320, 100
    seq = np.stack([synth.frame(h, w, 70), dot_lattice(h, w, 8), synth.frame(h, w, 71), synth.frame(h, w, 72)])
    ns = NativeSequence(w, h, nfeatures=nf, max_frames=T, chunk=T, host_chunk=2)
    out = ns.alloc_host_outputs()
    ns.run_host(torch.from_numpy(seq).pin_memory(), out)
    with pytest.raises(orbx.OrbxError) as e:
        ns.sync()
    assert e.value.code == -2
    n = out["n"].numpy()
    assert n[1] == ns.cap and n.max() <= ns.cap and n[0] > 0 and n[2] > 0
    ns.sync()                                                                  # reported once


def test_bench_configuration_1000_frames_vs_cv2():
    """The pass bench.py times - BASELINE configs[1]: the 1000-frame 1241x376 sequence, nfeatures 2000, ONE 1000-frame
    launch set on the device (chunk=1000, tensor-core matcher) and the 167-frame pipeline of orbx_sequence_run_host -
    compared with cv2 (the OpenCV code the reference binds) on 32 sampled frames and 32 sampled pairs, with the
    size-independent properties on all 999 pairs, and the two passes with each other on every byte."""
    import torch
    from oracle import synth, cv_ref, restated as R
    import slam_rs_b200 as orbx
    from slam_rs_b200.frontend import NativeSequence
    from tests.util import compare_orb
    T, h, w, nf = 1000, 376, 1241, 2000
    seq = synth.sequence(h, w, T)
    ns = NativeSequence(w, h, nfeatures=nf, max_frames=T, chunk=1000, host_chunk=167)
    cap = ns.cap
    dev = torch.from_numpy(seq).cuda()
    torch.cuda.synchronize()
    ns.run_device(dev)
    res = ns.results_to_host(T)
    n, ng = res["n"], res["n_good"]
    assert n.min() > 0 and n.max() <= cap and ng[1:].min() > 0
    kps = res["kps"].view(orbx.KEYPOINT_DTYPE).reshape(T, cap)
    good = res["good"].view(orbx.DMATCH_DTYPE).reshape(T, cap)
    rng = np.random.default_rng(3)
    frames = sorted(set([0, T - 1, 166, 167, 500] + list(rng.choice(T, 27, replace=False))))
    for t in frames:                                                 # extraction vs cv2, bit for bit
        st = compare_orb(kps[t, :n[t]], res["desc"][t, :n[t]], cv_ref.orb_detect_and_compute(seq[t], nf), h, w)
        assert st["missing"] == 0 and st["extra"] == 0 and st["desc_bits_diff"] == 0, (t, st)
        assert st["resp_eq"] == st["angle_eq"] == st["pt_eq"] == st["common"], (t, st)
    pairs = sorted(set([1, T - 1, 167, 168] + list(rng.choice(np.arange(1, T), 28, replace=False))))
    for t in pairs:                                                  # matching vs cv2 BFMatcher on the same rows
        d0, d1 = res["desc"][t - 1, :n[t - 1]], res["desc"][t, :n[t]]
        idx, dist, c = cv_ref.knn2(d0, d1)
        mutual = np.zeros(len(d0), bool)
        mutual[cv_ref.match_k1(d0, d1, cross_check=True)[:, 0]] = True
        keep = R.ratio_test(idx, dist, c, 0.75) & mutual
        g = good[t, :ng[t]]
        assert np.array_equal(g["query_idx"], np.nonzero(keep)[0]), t
        assert np.array_equal(g["train_idx"], idx[keep, 0]) and np.array_equal(g["distance"], dist[keep, 0].astype(np.float32))
    for t in range(1, T):                                            # properties of every pair
        g = good[t, :ng[t]]
        assert np.all(np.diff(g["query_idx"]) > 0) and len(np.unique(g["train_idx"])) == len(g)
        q, tr = res["desc"][t - 1][g["query_idx"]], res["desc"][t][g["train_idx"]]
        assert np.array_equal(np.unpackbits(q ^ tr, axis=1).sum(1), g["distance"].astype(np.int64))
    # the end-to-end pass (host buffers, 167-frame pipeline stages) returns the same bytes
    out = ns.alloc_host_outputs()
    ns.run_host(torch.from_numpy(seq).pin_memory(), out)
    ns.sync()
    assert np.array_equal(out["n"].numpy(), n) and np.array_equal(out["n_good"].numpy()[1:], ng[1:])
    ok, od, og = out["kps"].numpy(), out["desc"].numpy(), out["good"].numpy()
    for t in range(T):
        assert np.array_equal(ok[t, :n[t]], res["kps"][t, :n[t]]) and np.array_equal(od[t, :n[t]], res["desc"][t, :n[t]]), t
        if t:
            assert np.array_equal(og[t, :ng[t]], res["good"][t, :ng[t]]), t
