"""GPU: the allocation-free entry point orbx_knn_match2_static (csrc/match.cu; no counterpart in the reference)
against the existing entry point orbx_knn_match2.

Collected last on purpose.  Integer work: the tolerance is the one every matcher test
uses - bit-exact records and counts.  Two calls in a row share the static scratch array; the second must not see the
first one's contents.  Passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _static_knn(orbx, torch, q, t):
    dq, dt = torch.from_numpy(q).cuda(), torch.from_numpy(t).cuda()
    out = torch.full((max(len(q), 1), 2, 16), 0xAB, dtype=torch.uint8, device="cuda")
    cnt = torch.full((max(len(q), 1),), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    vp = C.c_void_p
    rc = orbx.lib.orbx_knn_match2_static(0, vp(dq.data_ptr()), len(q), vp(dt.data_ptr()) if len(t) else None, len(t),
                                         vp(out.data_ptr()), vp(cnt.data_ptr()), None)
    assert rc == 0, orbx.lib.orbx_last_error()
    torch.cuda.synchronize()
    rec = out.cpu().numpy().view(orbx.DMATCH_DTYPE).reshape(-1, 2)[:len(q)]
    return rec, cnt.cpu().numpy()[:len(q)]


def test_static_entry_matches_existing_entry_twice_in_a_row():
    import torch
    import slam_rs_b200 as orbx
    from oracle import synth
    ref = orbx.FeatureMatcher(mode="popc")
    launches0 = orbx.launch_count()
    # call 1: few queries, many train rows -> 30+ train ranges through the static scratch; call 2: other sizes, ragged,
    # fewer ranges (stale keys of call 1 stay in the array and must not be read); call 3: one range (no scratch at all)
    for nq, nt, seed in ((300, 100_001, 1), (1111, 33_333, 2), (5000, 200, 3)):
        q, t = synth.random_descriptors(nq, seed), synth.random_descriptors(nt, seed + 50)
        t[nt // 2] = t[nt - 1] = t[3]                      # ties across ranges: lowest train index wins
        q[0] = t[3]
        want, wcnt = ref.knn_match(q, t)
        got, gcnt = _static_knn(orbx, torch, q, t)
        assert np.array_equal(gcnt, wcnt)
        assert np.array_equal(got, want)
    assert orbx.launch_count() > launches0
    # empty sides (matcher.rs:29-31) and the stated limits
    rec, cnt = _static_knn(orbx, torch, synth.random_descriptors(9, 4), np.zeros((0, 32), np.uint8))
    assert np.all(cnt == 0)
    vp = C.c_void_p
    big = torch.zeros((8193, 32), dtype=torch.uint8, device="cuda")
    out = torch.zeros((8193, 2, 16), dtype=torch.uint8, device="cuda")
    cnt = torch.zeros((8193,), dtype=torch.int32, device="cuda")
    rc = orbx.lib.orbx_knn_match2_static(0, vp(big.data_ptr()), 8193, vp(big.data_ptr()), 100, vp(out.data_ptr()),
                                         vp(cnt.data_ptr()), None)
    assert rc == -2 and b"8192" in orbx.lib.orbx_last_error()
    host = np.zeros((4, 32), np.uint8)
    rc = orbx.lib.orbx_knn_match2_static(0, vp(host.ctypes.data), 4, vp(big.data_ptr()), 100, vp(out.data_ptr()),
                                         vp(cnt.data_ptr()), None)
    assert rc == -1                                         # host pointers would need staging = an allocation
