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


def test_native_bad_arguments():
    import slam_rs_b200 as s
    from slam_rs_b200.frontend import NativeSequence
    with pytest.raises(s.OrbxError):
        NativeSequence(8, 8, max_frames=4)                                     # below the 16 px minimum
    ns = NativeSequence(320, 240, nfeatures=100, max_frames=4, chunk=4, host_chunk=2)
    import torch
    with pytest.raises(s.OrbxError):
        ns.run_device(torch.zeros((5, 240, 320), dtype=torch.uint8, device="cuda"))   # T > max_frames
