"""GPU parity: the multi-GPU entry of the C ABI (orbx_comm_*, csrc/comm.cu) - query-sharded map-to-frame matching
(/root/reference/src/mapping/map.rs:107 at BASELINE configs[3] sizes) - against the single-GPU matcher and the oracle.

A one-GPU box runs the split as two / three VIRTUAL ranks on the same device (same slicing, same global query_idx, the
same peer-pointer epilogue stores; NCCL is skipped); a box with >= 2 GPUs also runs real ranks with the NCCL broadcast."""
import numpy as np
import pytest

from oracle import restated as R, synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def orbx():
    import slam_rs_b200 as s
    return s


def device_lists():
    import torch
    n = torch.cuda.device_count()
    lists = [[0], [0, 0], [0, 0, 0]]
    if n >= 2:
        lists.append([0, 1])
    if n >= 4:
        lists.append([0, 1, 2, 3])
    if n >= 8:
        lists.append(list(range(8)))
    return lists


def test_sharded_knn_equals_single_gpu_and_oracle(orbx):
    from slam_rs_b200.mapmatch import ShardedMatcher
    m, f = synth.map_and_frame_descriptors(20000, 1003, f=0)   # odd query count: uneven slices
    ri, rd, rc = R.knn(f, m, 2)
    single = orbx.FeatureMatcher()
    k_ref, c_ref = single.knn_match(f, m)
    assert np.array_equal(k_ref["train_idx"], ri) and np.array_equal(c_ref, rc)
    for devs in device_lists():
        sm = ShardedMatcher(devs)
        assert sm.size() == len(devs)
        sm.set_train(m)
        knn, cnt = sm.knn_match(f)
        assert np.array_equal(knn, k_ref), devs          # every field, query_idx global
        assert np.array_equal(cnt, c_ref), devs
        k1 = sm.match_descriptors(f)
        assert np.array_equal(k1["train_idx"], ri[:, 0]) and np.array_equal(k1["query_idx"], np.arange(len(f)))
        assert np.array_equal(k1["distance"], rd[:, 0].astype(np.float32))
        # fewer query rows than ranks x 8, and the empty guards of matcher.rs:29-31
        few, cf = sm.knn_match(f[:5])
        assert np.array_equal(few, k_ref[:5]) and np.array_equal(cf, c_ref[:5])
        assert len(sm.match_descriptors(f[:0])) == 0
        sm.set_train(m[:0])
        assert len(sm.match_descriptors(f)) == 0
        e, ce = sm.knn_match(f[:7])
        assert np.all(ce == 0)
        sm.close()


def test_sharded_device_pointers_and_train_update(orbx):
    """Device-resident queries / results on devices[0] (asynchronous path) and a second set_train (map update)."""
    import torch
    from slam_rs_b200.mapmatch import ShardedMatcher
    m, f = synth.map_and_frame_descriptors(40000, 640, f=1)
    for devs in device_lists()[1:]:
        sm = ShardedMatcher(devs)
        q = torch.from_numpy(f).cuda(0)
        out = torch.zeros((len(f), 2, 16), dtype=torch.uint8, device="cuda:0")
        cnt = torch.zeros((len(f),), dtype=torch.int32, device="cuda:0")
        torch.cuda.synchronize()
        # the larger sets re-allocate and re-broadcast; from 16384 rows on the ranks keep an expanded copy of the train set
        # for the tensor-core kernel, which every set_train must refresh
        for train in (m[:4000], m[:20000], m[20000:], m):
            sm.set_train(torch.from_numpy(train).cuda(0))
            sm.knn_match_device(q, out, cnt)
            sm.sync()
            ri, rd, rc = R.knn(f, train, 2)
            rec = out.cpu().numpy().view(orbx.DMATCH_DTYPE).reshape(len(f), 2)
            assert np.array_equal(rec["train_idx"], ri) and np.array_equal(rec["distance"], rd.astype(np.float32)), devs
            assert np.array_equal(rec["query_idx"][:, 0], np.arange(len(f))) and np.array_equal(cnt.cpu().numpy(), rc)
        sm.set_timing(True)
        sm.knn_match_device(q, out, cnt)
        ms, calls, k0 = sm.timing()
        assert calls == 1 and ms > 0 and k0 > 0
        sm.close()
