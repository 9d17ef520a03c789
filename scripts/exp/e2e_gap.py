"""Where do the 0.7 ms between the device-resident pass (11.5 ms) and the end-to-end pass (12.3 ms) go?
(a) run_host with every output, (b) without any download (NULL outputs), (c) the copies alone, both directions at once."""
import ctypes as C
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
import bench  # noqa: E402
from slam_rs_b200.frontend import NativeSequence  # noqa: E402

synth = bench.load_synth()
T, H, W = 1000, 376, 1241
frames = torch.from_numpy(synth.sequence(H, W, T, canvas_seed=100, noise_seed0=1000)).pin_memory()
for hc in (250,):
    ns = NativeSequence(W, H, nfeatures=2000, max_frames=T, chunk=1000, host_chunk=hc)
    outs = [ns.alloc_host_outputs(), ns.alloc_host_outputs()]
    vp = C.c_void_p

    def run(i, download):
        o = outs[i & 1]
        if download == "all":
            ns.run_host(frames, o)
        else:
            args = [vp(o[k].data_ptr()) if k in download else None for k in ("kps", "desc", "n", "good", "n_good")]
            ns._check(ns._lib.orbx_sequence_run_host(ns._h, vp(frames.data_ptr()), T, *args))

    for name, dl in (("all outputs", "all"), ("no download", ()), ("counts only", ("n", "n_good")),
                     ("kps+desc only", ("kps", "desc", "n")), ("good only", ("good", "n_good"))):
        for i in range(4):
            run(i, dl)
        ns.sync()
        ns.set_timing(True)
        for i in range(20):
            run(i, dl)
        ns.sync()
        ms, p = ns.elapsed()
        ns.set_timing(False)
        print("host_chunk %d  %-14s %.3f ms per pass" % (hc, name, ms / p), flush=True)
    dev = frames.cuda()
    for i in range(3):
        ns.run_device(dev)
    ns.sync()
    ns.set_timing(True)
    for i in range(20):
        ns.run_device(dev)
    ns.sync()
    ms, p = ns.elapsed()
    print("host_chunk %d  %-14s %.3f ms per pass" % (hc, "run_device", ms / p), flush=True)
    ns.close()

# the copies alone
dev = torch.empty_like(frames, device="cuda")
res_dev = torch.empty((178_000_000,), dtype=torch.uint8, device="cuda")
res_host = torch.empty((178_000_000,), dtype=torch.uint8).pin_memory()
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for mode in ("h2d", "d2h", "both"):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        if mode in ("h2d", "both"):
            with torch.cuda.stream(s1):
                dev.copy_(frames, non_blocking=True)
        if mode in ("d2h", "both"):
            with torch.cuda.stream(s2):
                res_host.copy_(res_dev, non_blocking=True)
    torch.cuda.synchronize()
    print("copies alone, %-5s %.3f ms per pass (467 MB up, 178 MB down)" % (mode, (time.perf_counter() - t0) * 100), flush=True)
