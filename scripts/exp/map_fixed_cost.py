"""Fixed cost of one map-to-frame step: the per-rank problems of an N-GPU split (5000 / N queries x 1 M train rows) timed on ONE
GPU through ShardedMatcher([0]) - CUDA events around the whole launch set (orbx_comm timing) and the ncu-free kernel split by
event pairs is not available, so the launch list of this script under ncu gives the per-kernel times."""
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import bench  # noqa: E402
from slam_rs_b200.mapmatch import ShardedMatcher  # noqa: E402

synth = bench.load_synth()
m, _ = synth.map_and_frame_descriptors(1_000_000, 5000, f=0)
rng = np.random.default_rng(8)
sm = ShardedMatcher([0])
sm.set_train(m)
for nq in (5000, 2500, 1250, 625, 313, 128):
    q = torch.from_numpy(rng.integers(0, 256, (nq, 32), dtype=np.uint8)).cuda(0)
    out = torch.zeros((nq, 2, 16), dtype=torch.uint8, device="cuda:0")
    cnt = torch.zeros((nq,), dtype=torch.int32, device="cuda:0")
    for _ in range(3):
        sm.knn_match_device(q, out, cnt)
    sm.sync()
    sm.set_timing(True)
    t0 = time.perf_counter()
    for _ in range(20):
        sm.knn_match_device(q, out, cnt)
    sm.sync()
    host = (time.perf_counter() - t0) / 20 * 1e3
    t = sm.timing()
    sm.set_timing(False)
    print("nq %5d: %s host %.3f ms per step; ideal at 4.83 T pairs/s: %.3f ms" % (nq, t, host, nq * 1e6 / 4.83e12 * 1e3), flush=True)
