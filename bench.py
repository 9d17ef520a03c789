#!/usr/bin/env python
"""bench.py — ORB extract + frame-to-frame Hamming match throughput (frames/s @1241x376).

Workload (BASELINE.json configs[1]): a KITTI-shaped synthetic sequence of T=1000
grayscale 1241x376 frames per GPU, nfeatures=2000.  One step = one pass of the hot path over the
sequence: S1-S9 on every frame and the T-1 (t-1 -> t) matching problems (BFMatcher k=2 + cross-check
mutual test + Lowe ratio 0.75).  N>1: every rank owns an independent sequence (weak scaling, no
data-path collective — SURVEY.md §8e); value = all frames / max-over-ranks time.

A step is `--passes` (10) passes over the resident sequence, so that a --steps 20 run times > 2 s of GPU work.
`value`: frames resident in HBM, timed with CUDA events.  `e2e`: the same passes through the reference-facing plugin
call orbx_sequence_run_host (csrc/sequence.cu) with HOST buffers - pinned frames in, keypoints / descriptors / good
matches back out - timed by the library's own CUDA events.  Attached sub-measurements (`--subs 1`): `hd` (configs[2]
shape), `seqs` (configs[4] shape: 8 pre-staged sequences per GPU), `map` (configs[3]: orbx_comm over all GPUs of the
run, driven by rank 0), `latency` (configs[0]: two frames, per-call).

  python bench.py [--gpus N] [--steps K] [--warmup W]          our arm (liborbx on B200)
  python bench.py --impl reference ...                          reference arm: cv2 (= the OpenCV code the
                                                                reference binds) on the host cores
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W_IMG, H_IMG, NFEAT, T_SEQ = 1241, 376, 2000, 1000
RATIO = 0.75
METRIC = "ORB extract+match frames/s @1241x376"
WORKLOAD = ("C2 KITTI-shaped synthetic sequence 1241x376 x %d frames/GPU, nfeatures=2000, 8 levels, "
            "frame-to-frame BFMatcher Hamming k=2 + cross-check + Lowe 0.75")


def make_config(args):
    """The `config` object of the JSON line: the same keys and values in both arms (ours and --impl reference)."""
    return {"workload": WORKLOAD % args.frames, "frames_per_gpu": args.frames, "width": W_IMG, "height": H_IMG,
            "nfeatures": NFEAT, "levels": 8, "matching": "k=2 + cross-check + Lowe %.2f" % RATIO,
            "passes_per_step": args.passes,
            "step": "%d passes over the resident %d-frame sequence (ours); a %d-frame prefix of the same sequence "
                    "(reference arm, bounded CPU sample)" % (args.passes, args.frames, args.ref_frames),
            "l2": "inputs larger than L2 (%.0f MB of frames per pass)" % (args.frames * H_IMG * W_IMG / 1e6)}


def load_synth():
    import importlib.util as u
    spec = u.spec_from_file_location("_orbx_synth", os.path.join(ROOT, "slam-rs_b200", "synth.py"))
    m = u.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            j = json.load(open(p))
            return float(j["hbm_gbs"]), "measured", j
        except Exception:
            pass
    return 6650.0, "fallback", {}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        # NVML from a thread (one sample every ~2 ms: the timed region is tens of ms); nvidia-smi -lms as fallback
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml = pynvml
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.samples, self.mask, self.running = [], 0, True
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _poll(self):
        n = self.nvml
        while self.running:
            try:
                self.samples.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
                self.mask |= int(n.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
            except Exception:
                pass
            time.sleep(0.002)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def _stop_nvml(self):
        self.running = False
        self.thread.join(timeout=1)
        bits = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap"}
        sm = sorted(self.samples)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": self.max_mhz, "samples": len(sm),
                "reasons": sorted(v for k, v in bits.items() if self.mask & k), "source": "nvml"}

    def stop(self):
        if self.nvml:
            return self._stop_nvml()
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------
# CPU reference: cv2 == the cv::ORB / cv::BFMatcher code the reference calls (detector.rs:32, matcher.rs:34)
# ---------------------------------------------------------------------------------------------------
def cpu_pass(frames, threads, reference_filter=False):
    """One pass of the same hot path with OpenCV on the host: returns (seconds, frames_done).  The match lists are
    built (as the reference builds its Vector<DMatch>) but the Python-side filtering of the k=2 + cross-check + Lowe
    composition is kept outside the clock: OpenCV cannot express it in one call, so the reference arm pays for its two
    BFMatcher passes only.  reference_filter: the reference's own loop instead (k=1 + filter_good_matches(2.0))."""
    import cv2
    cv2.setNumThreads(threads)
    orb = cv2.ORB_create(nfeatures=500 if reference_filter else NFEAT)
    bf = cv2.BFMatcher(cv2.NORM_HAMMING, False)
    bfx = cv2.BFMatcher(cv2.NORM_HAMMING, True)
    t_total = 0.0
    prev = None
    for img in frames:
        t0 = time.perf_counter()
        _, desc = orb.detectAndCompute(img, None)
        knn = mutual = m1 = None
        if prev is not None and desc is not None and len(prev) and len(desc):
            if reference_filter:
                m1 = bf.match(prev, desc)
            else:
                knn = bf.knnMatch(prev, desc, k=2)
                mutual = bfx.match(prev, desc)
        t_total += time.perf_counter() - t0
        if knn is not None:   # outside the clock: Lowe + mutual composition in Python
            ok = {m.queryIdx for m in mutual}
            _ = [p[0] for p in knn if len(p) == 2 and p[0].queryIdx in ok
                 and np.float32(p[0].distance) < np.float32(RATIO) * np.float32(p[1].distance)]
        prev = desc
    return t_total, len(frames)


def cpu_baseline(frames, threads, budget_s=12.0):
    """cv2 on the host cores on a bounded prefix of the same sequence: all cores and one thread."""
    import cv2
    cpu_pass(frames[:3], threads)  # warm-up
    dt, n = cpu_pass(frames[:min(len(frames), 400)], threads)
    n1 = max(8, min(len(frames), int(budget_s * (n / dt) / max(threads, 1) * 1.5)))
    dt1, n1 = cpu_pass(frames[:n1], 1)
    dtr, nr = cpu_pass(frames[:min(len(frames), 200)], threads, reference_filter=True)
    return {"value": n / dt, "unit": "frames/s", "cores": threads, "kind": "reference",
            "sample": "cv2 %s ORB_create(nfeatures=2000).detectAndCompute + BFMatcher knnMatch(k=2) + crossCheck match on "
                      "the first %d frames of the same sequence, %d OpenCV threads, %.1f s (Python-side Lowe / mutual "
                      "filtering outside the clock)" % (cv2.__version__, n, threads, dt),
            "one_thread": {"value": n1 / dt1, "unit": "frames/s", "cores": 1,
                           "sample": "same, cv2.setNumThreads(1), first %d frames, %.1f s" % (n1, dt1)},
            "reference_faithful": {"value": nr / dtr, "unit": "frames/s", "cores": threads,
                                   "sample": "the reference's own loop: ORB::create_def() (500 features) + match (k=1) + "
                                             "filter_good_matches(2.0) source, first %d frames, %.1f s" % (nr, dtr)}}


def run_reference(args, rank, world):
    if rank != 0:
        return
    synth = load_synth()
    threads = os.cpu_count() or 1
    nsample = args.ref_frames
    frames = synth.sequence(H_IMG, W_IMG, nsample)
    import cv2
    for _ in range(args.warmup):
        cpu_pass(frames[:4], threads)
    times = []
    for _ in range(args.steps):
        dt, n = cpu_pass(frames, threads)
        times.append(dt)
    tot = sum(times)
    val = nsample * args.steps / tot
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * tot / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": make_config(args),
        "cpu_baseline": {"value": val, "unit": "frames/s", "cores": threads, "kind": "reference",
                         "sample": "cv2 %s (the OpenCV code the reference binds) on the first %d frames of the sequence "
                                   "per step, %d threads" % (cv2.__version__, nsample, threads)},
        "e2e": {"value": val, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ---------------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------------
def pyramid_pixels(h, w):
    sf = np.float32(1.2)
    tot = []
    for l in range(8):
        inv = np.float32(1.0) / np.float32(np.power(np.float64(sf), np.float64(l)))
        tot.append(int(np.rint(np.float32(h) * inv)) * int(np.rint(np.float32(w) * inv)))
    return tot


def fast_tiles(h, w):
    """FAST tiles (62 x 30 interior of the 31-px-culled level) per frame — mirrors build_geom in detector_run.inc."""
    n = 0
    sf = np.float32(1.2)
    for l in range(8):
        inv = np.float32(1.0) / np.float32(np.power(np.float64(sf), np.float64(l)))
        lh, lw = int(np.rint(np.float32(h) * inv)), int(np.rint(np.float32(w) * inv))
        iw, ih = lw - 62, lh - 62
        if iw > 0 and ih > 0:
            n += -(-iw // 62) * -(-ih // 30)
    return n


def sub_hd(args, torch, dist, world, local_rank, synth, barrier):
    """configs[2] shape: 1920x1080 frames, nfeatures 5000, per-GPU batch (sequence-sharded like the main workload)."""
    from slam_rs_b200.frontend import NativeSequence
    T, w, h, nf = args.hd_frames, 1920, 1080, 5000
    base = np.stack([synth.frame(h, w, 200 + i + 100 * local_rank) for i in range(min(T, 12))])
    frames = np.concatenate([base] * (-(-T // len(base))))[:T].copy()
    for i in range(T):  # distinct frames: a per-frame brightness offset keeps the repeated textures from being identical
        frames[i] = np.clip(frames[i].astype(np.int16) + (i // len(base)) * 3, 0, 255).astype(np.uint8)
    ns = NativeSequence(w, h, nfeatures=nf, max_frames=T, chunk=T, host_chunk=max(T // 4, 1), device=local_rank,
                        ratio=RATIO, cross_check=True)
    pinned = torch.from_numpy(frames).pin_memory()
    dev = torch.from_numpy(frames).cuda(local_rank)
    out = ns.alloc_host_outputs()
    res = {}
    for name, run in (("value", lambda: ns.run_device(dev)), ("e2e", lambda: ns.run_host(pinned, out))):
        for _ in range(3):
            run()
        ns.sync()
        barrier()
        ns.set_timing(True)
        for _ in range(args.hd_passes):
            run()
        ns.sync()
        ms, passes = ns.elapsed()
        ns.set_timing(False)
        t = torch.tensor([ms], dtype=torch.float64, device="cuda:%d" % local_rank)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res[name] = T * world * passes / (float(t[0]) * 1e-3)
        res[name + "_ms_per_pass"] = float(t[0]) / passes
    n = out["n"][:T].numpy()
    ns.close()
    return {"workload": "C3 shape: 1920x1080 synthetic frames, nfeatures=5000, %d frames per GPU per pass, frame-to-frame "
                        "k=2 + cross-check + Lowe" % T, "n_gpus": world, "value": res["value"], "unit": "frames/s",
            "ms_per_pass": res["value_ms_per_pass"], "e2e": {"value": res["e2e"], "ms_per_pass": res["e2e_ms_per_pass"],
                                                             "api": "orbx_sequence_run_host"},
            "keypoints_per_frame": float(n.mean()), "l2": "inputs larger than L2 (%.0f MB per pass)" % (T * w * h / 1e6)}


def sub_seqs(args, torch, dist, world, local_rank, frames_pinned, barrier):
    """configs[4]: 64 independent sequences on 8 GPUs = S = 8 sequences of 1000 frames per GPU (29.9 GB of frames over the
    box), all pre-staged - in HBM for `value`, in pinned host memory for `e2e` - and run back to back through the native
    sequence front end.  Sequences 1..7 of a rank are derived from its sequence 0 on the device (horizontal / vertical
    flip / time reversal, the 8 combinations): different images, keypoints and descriptors, no extra host synthesis."""
    from slam_rs_b200.frontend import NativeSequence
    S = args.seqs
    T = frames_pinned.shape[0]
    base = frames_pinned.cuda(local_rank, non_blocking=True)
    dev_seqs, host_seqs = [], []
    for k in range(S):
        v = base
        dims = [d for b, d in ((1, 2), (2, 1), (4, 0)) if k & b]
        if dims:
            v = torch.flip(base, dims)
        v = v.contiguous()
        dev_seqs.append(v)
        h = torch.empty(v.shape, dtype=v.dtype).pin_memory()
        h.copy_(v)
        host_seqs.append(h)
    torch.cuda.synchronize()
    ns = NativeSequence(W_IMG, H_IMG, nfeatures=NFEAT, max_frames=T, chunk=args.chunk, host_chunk=args.host_chunk,
                        device=local_rank, ratio=RATIO, cross_check=True)
    outs = [ns.alloc_host_outputs(), ns.alloc_host_outputs()]
    res, kp = {}, []
    for name, run in (("value", lambda k: ns.run_device(dev_seqs[k])),
                      ("e2e", lambda k: ns.run_host(host_seqs[k], outs[k & 1]))):
        for k in range(min(S, 2)):
            run(k)
        ns.sync()
        barrier()
        ns.set_timing(True)
        for k in range(S):
            run(k)
            if name == "e2e" and k >= 1:
                ns.wait_pass(1)  # sequence k-1 is complete on the host while sequence k runs: consume it
                kp.append(float(outs[(k - 1) & 1]["n"][:T].numpy().mean()))
        ns.sync()
        if name == "e2e":
            kp.append(float(outs[(S - 1) & 1]["n"][:T].numpy().mean()))
        ms, passes = ns.elapsed()
        ns.set_timing(False)
        t = torch.tensor([ms], dtype=torch.float64, device="cuda:%d" % local_rank)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res[name] = T * world * passes / (float(t[0]) * 1e-3)
        res[name + "_ms"] = float(t[0]) / passes
    ns.close()
    return {"workload": "C5 shape: %d independent 1241x376 sequences of %d frames (%d per GPU), pre-staged, frame-to-frame "
                        "k=2 + cross-check + Lowe" % (S * world, T, S), "n_gpus": world, "sequences": S * world,
            "staged_gb": S * world * T * H_IMG * W_IMG / 1e9, "value": res["value"], "unit": "frames/s",
            "ms_per_sequence": res["value_ms"],
            "e2e": {"value": res["e2e"], "ms_per_sequence": res["e2e_ms"], "api": "orbx_sequence_run_host"},
            "keypoints_per_frame_by_sequence": kp}


def sub_map(args, torch, world, synth):
    """configs[3]: 1 M map descriptors (train, resident on every GPU, one NCCL broadcast) vs 5000 frame descriptors
    (query, sharded over the GPUs), k = 2 - through the C ABI's multi-GPU entry (orbx_comm_*): ONE process (this rank 0)
    drives all `world` GPUs, the other ranks of the bench wait.  Strong scaling: 5e9 pairs per step whatever N."""
    from slam_rs_b200.mapmatch import ShardedMatcher
    rows, nq, nframes = 1_000_000, 5000, 4
    m, _ = synth.map_and_frame_descriptors(rows, nq, f=0)
    rng = np.random.default_rng(8)
    frames = []
    for f in range(nframes):
        pick = rng.choice(rows, nq // 2, replace=False)
        flips = rng.random((nq // 2, 256)) < 0.1
        noisy = m[pick] ^ np.packbits(flips, axis=1, bitorder="little")
        frames.append(np.concatenate([noisy, rng.integers(0, 256, (nq - nq // 2, 32), dtype=np.uint8)]))
    sm = ShardedMatcher(list(range(world)))
    t0 = time.perf_counter()
    sm.set_train(m)
    bcast_ms = (time.perf_counter() - t0) * 1e3
    q = [torch.from_numpy(f).cuda(0) for f in frames]
    out = torch.zeros((nq, 2, 16), dtype=torch.uint8, device="cuda:0")
    cnt = torch.zeros((nq,), dtype=torch.int32, device="cuda:0")
    torch.cuda.synchronize()
    for i in range(3):
        sm.knn_match_device(q[i % nframes], out, cnt)
    sm.sync()
    sm.set_timing(True)
    t0 = time.perf_counter()
    for i in range(args.map_steps):
        sm.knn_match_device(q[i % nframes], out, cnt)
    sm.sync()
    wall_ms = (time.perf_counter() - t0) * 1e3
    ms, calls, k0 = sm.timing()
    # end to end: query descriptors in host memory, records back in host memory
    sm.set_timing(False)
    t0 = time.perf_counter()
    for i in range(args.map_steps):
        knn, c = sm.knn_match(frames[i % nframes])
    e2e_ms = (time.perf_counter() - t0) * 1e3
    sm.close()
    step = ms / max(calls, 1)
    return {"workload": "C4 map-to-frame: 1M map descriptors resident on every GPU (one ncclBroadcast per map update) vs "
                        "5000 frame descriptors, query rows sharded over the GPUs, k=2; records stored straight into GPU 0's "
                        "buffer through peer-mapped pointers", "api": "orbx_comm_knn_match2 (single process, rank 0)",
            "n_gpus": world, "pairs_per_s": float(rows) * nq / (step * 1e-3), "unit": "pairs/s", "scaling": "strong",
            "ms_per_step": step, "kernel_ms": k0 / max(calls, 1), "comm_ms": max(step - k0 / max(calls, 1), 0.0),
            "host_ms_per_step": wall_ms / args.map_steps, "train_broadcast_ms": bcast_ms,
            "e2e": {"pairs_per_s": float(rows) * nq / (e2e_ms / args.map_steps * 1e-3), "ms_per_step": e2e_ms / args.map_steps,
                    "h2d_bytes_per_step": nq * 32, "d2h_bytes_per_step": nq * (2 * 16 + 4)},
            "steps": args.map_steps, "timer": "CUDA events on GPU 0's stream: queries staged -> every rank's records landed"}


def sub_latency(args, orbx, synth, local_rank):
    """configs[0]: two 1241x376 frames, per-call latency through the per-frame API with HOST buffers - what every
    reference caller does (examples/visual_odometry.rs:102-107, one frame at a time) - beside cv2 on the host."""
    import cv2
    a, b = synth.frame(H_IMG, W_IMG, 0), synth.frame(H_IMG, W_IMG, 1)
    det = orbx.OrbDetector(NFEAT, device=local_rank, max_width=W_IMG, max_height=H_IMG)
    m = orbx.FeatureMatcher(device=local_rank, cross_check=True)

    def ours():
        _, d1 = det.detect_and_compute(a)
        _, d2 = det.detect_and_compute(b)
        knn, cnt = m.knn_match(d1, d2)
        return m.ratio_test(knn, cnt, RATIO)

    def timed(fn, n):
        ts = []
        for _ in range(n):
            t0 = time.perf_counter()
            fn()
            ts.append((time.perf_counter() - t0) * 1e6)
        ts = np.sort(np.array(ts))
        return float(ts[len(ts) // 2]), float(ts[min(len(ts) - 1, int(len(ts) * 0.99))])

    res = {}
    for graph in (0, 1):
        try:
            det.set_graph(bool(graph))
        except Exception:
            if graph:
                continue
        for _ in range(20):
            ours()
        p50, p99 = timed(ours, args.latency_iters)
        res["graph" if graph else "stream"] = {"p50_us": p50, "p99_us": p99}
    good = ours()
    threads = os.cpu_count() or 1
    cv2.setNumThreads(threads)
    orb = cv2.ORB_create(nfeatures=NFEAT)
    bf, bfx = cv2.BFMatcher(cv2.NORM_HAMMING, False), cv2.BFMatcher(cv2.NORM_HAMMING, True)

    def ref():
        _, d1 = orb.detectAndCompute(a, None)
        _, d2 = orb.detectAndCompute(b, None)
        bf.knnMatch(d1, d2, k=2)
        bfx.match(d1, d2)

    for _ in range(3):
        ref()
    c50, c99 = timed(ref, 20)
    return {"workload": "C1: ORB nfeatures=2000 on two 1241x376 frames + BFMatcher k=2 + cross-check + Lowe, one call per "
                        "frame, host buffers in and out", "unit": "us per frame pair",
            "ours": res, "good_matches": int(len(good)),
            "cv2": {"p50_us": c50, "p99_us": c99, "cores": threads}, "iters": args.latency_iters}


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — liborbx has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    host_group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        # host-side barrier for the phases in which rank 0 alone drives ALL GPUs (sub_map): a NCCL barrier would leave a
        # spinning kernel of every other rank on the GPUs rank 0 is timing (and time-slice against it)
        host_group = dist.new_group(backend="gloo")
    import slam_rs_b200 as orbx
    from slam_rs_b200.frontend import SequenceFrontEnd, NativeSequence
    synth = load_synth()

    T, P = args.frames, args.passes
    if world > 1 and args.numa_bind:
        bind_near_gpu(local_rank)  # before the pinned buffers are first touched
    # every rank owns an independent sequence (canvas seed differs): weak scaling, no collective
    frames = synth.sequence(H_IMG, W_IMG, T, canvas_seed=100 + rank, noise_seed0=1000 + 100000 * rank)
    fe = SequenceFrontEnd(W_IMG, H_IMG, nfeatures=NFEAT, max_frames=T, chunk=args.chunk, device=local_rank,
                          ratio=RATIO, cross_check=True, overlap=args.overlap,
                          host_chunk=args.host_chunk, dual=bool(args.dual and args.torch_e2e),
                          pose=orbx.CameraIntrinsics.kitti() if args.pose else None)
    if args.matcher != "auto":
        fe.matcher.set_mode(args.matcher)
    frames_pinned = torch.from_numpy(frames).pin_memory()
    frames_dev = torch.zeros((T, H_IMG, fe.pitch), dtype=torch.uint8, device=dev)
    frames_dev[:, :, :W_IMG].copy_(frames_pinned)
    torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident: warm-up, then K timed steps of P passes with per-stage event timers on
    for _ in range(args.warmup):
        for _ in range(P):
            fe.run_device(frames_dev)
    barrier()
    inline_timers = bool(args.stage_timers)
    if inline_timers:
        fe.det.set_timing(True)
        fe.matcher.set_timing(True)
        fe.det.timing(reset=True)
        fe.matcher.timing(reset=True)
    launches0 = orbx.launch_count()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps * P):
        fe.run_device(frames_dev)
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    launches = orbx.launch_count() - launches0
    stage_passes = args.steps * P
    if not inline_timers:
        # --stage-timers 0: the timed region above ran without the per-stage events; the stage breakdown comes from P
        # further passes with the timers on (measured difference of the headline: +0.3 %)
        fe.det.set_timing(True)
        fe.matcher.set_timing(True)
        fe.det.timing(reset=True)
        fe.matcher.timing(reset=True)
        for _ in range(P):
            fe.run_device(frames_dev)
        torch.cuda.synchronize()
        stage_passes = P
    stage = fe.det.timing(reset=True)
    match_ms, match_launches = fe.matcher.timing(reset=True)
    fe.det.set_timing(False)
    fe.matcher.set_timing(False)
    n_dev = np.minimum(fe.n[:T].cpu().numpy().astype(np.int64), fe.cap)
    n_good_dev = int(fe.n_good[1:T].cpu().numpy().astype(np.int64).sum())

    # ---- end to end through the reference-facing plugin call with HOST buffers: orbx_sequence_run_host (csrc/
    # sequence.cu - the entry a Rust / C++ host binds for the loop of examples/visual_odometry.rs:84-218): pinned host
    # frames in, keypoints / descriptors / good matches back in pinned host arrays, every pass.  Timed on the device by
    # the library's own CUDA events (orbx_sequence_elapsed): first upload -> last download.
    ns = NativeSequence(W_IMG, H_IMG, nfeatures=NFEAT, max_frames=T, chunk=args.chunk, host_chunk=args.host_chunk,
                        device=local_rank, ratio=RATIO, cross_check=True)
    # consecutive run_host calls are pipelined into each other (sequence.cu): as a streaming host would, the passes
    # alternate two sets of pinned result arrays, so the set of pass k stays untouched while pass k+1 runs
    host_outs = [ns.alloc_host_outputs(), ns.alloc_host_outputs()]
    host_out = host_outs[(args.steps * P - 1) & 1]  # the set the last timed pass writes
    for i in range(max(2, args.warmup)):
        ns.run_host(frames_pinned, host_outs[i & 1])
    ns.sync()
    barrier()
    ns.set_timing(True)
    t0 = time.perf_counter()
    for i in range(args.steps * P):
        ns.run_host(frames_pinned, host_outs[i & 1])
    ns.sync()
    e2e_wall_ms = (time.perf_counter() - t0) * 1e3
    ms_e2e, e2e_passes = ns.elapsed()
    ns.set_timing(False)
    barrier()
    clocks = sampler.stop()
    n_good = int(host_out["n_good"][1:T].numpy().astype(np.int64).sum())
    e2e_cap = ns.cap
    ns.close()
    # H2D / D2H rate of this rank's host link, copying alone (the end-to-end pass at 8 ranks is bound by the box's
    # aggregate host bandwidth - DESIGN.md §4 - so the line carries the measured rates)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    stage_dev = torch.empty_like(frames_pinned, device=dev)
    back = torch.empty_like(frames_pinned).pin_memory()
    stage_dev.copy_(frames_pinned, non_blocking=True)
    barrier()
    ev[0].record()
    stage_dev.copy_(frames_pinned, non_blocking=True)
    ev[1].record()
    back.copy_(stage_dev, non_blocking=True)
    ev[2].record()
    torch.cuda.synchronize()
    h2d_gbs = frames_pinned.numel() / (ev[0].elapsed_time(ev[1]) * 1e-3) / 1e9
    d2h_gbs = frames_pinned.numel() / (ev[1].elapsed_time(ev[2]) * 1e-3) / 1e9
    del stage_dev, back

    # ---- secondary: the same pass scheduled from Python over torch streams (round 1's e2e path)
    torch_e2e = None
    if args.torch_e2e:
        t_out = fe.alloc_host_outputs()
        for _ in range(2):
            fe.run_host(frames_pinned, t_out)
        barrier()
        e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e2.record()
        for _ in range(args.steps):
            fe.run_host(frames_pinned, t_out)
        e3.record()
        barrier()
        torch_e2e = e2.elapsed_time(e3) / args.steps

    cand0 = 0
    try:
        fe.det.detect_and_compute_device(frames_dev[:1], fe.kps[:1], fe.desc[:1], fe.n[:1], width=W_IMG)
        cand0 = sum(len(fe.det.candidates(l, 0)[0]) for l in range(8))
    except Exception:
        pass

    t = torch.tensor([ms_total, ms_e2e, torch_e2e or 0.0, -h2d_gbs, -d2h_gbs], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e = float(t[0]), float(t[1])
    torch_e2e = float(t[2]) if torch_e2e else None
    h2d_min, d2h_min = -float(t[3]), -float(t[4])
    passes_all = args.steps * P
    frames_all = T * world * passes_all
    value = frames_all / (ms_total * 1e-3)
    e2e_value = frames_all / (ms_e2e * 1e-3)

    # ---- sub-measurements the headline does not cover (north_star configs[0], [2], [3]); never fatal
    subs = {}
    del fe
    torch.cuda.empty_cache()
    if args.subs:
        try:
            hd = sub_hd(args, torch, dist, world, local_rank, synth, barrier)
            if rank == 0:
                subs["hd"] = hd
        except Exception as ex:  # noqa: BLE001
            subs["hd"] = {"error": repr(ex)[:300]}
        barrier()
        try:
            sq = sub_seqs(args, torch, dist, world, local_rank, frames_pinned, barrier)
            if rank == 0:
                subs["seqs"] = sq
        except Exception as ex:  # noqa: BLE001
            subs["seqs"] = {"error": repr(ex)[:300]}
        torch.cuda.empty_cache()
        barrier()
        if rank == 0:
            for name, fn in (("map", lambda: sub_map(args, torch, world, synth)),
                             ("latency", lambda: sub_latency(args, orbx, synth, local_rank))):
                try:
                    subs[name] = fn()
                except Exception as ex:  # noqa: BLE001
                    subs[name] = {"error": repr(ex)[:300]}
        torch.cuda.synchronize()
        if host_group is not None:
            dist.barrier(group=host_group)  # the other ranks wait on the host: their GPUs stay idle for rank 0

    if rank == 0:
        hbm_peak, peak_kind, peaks = measured_peaks()
        P_px = sum(pyramid_pixels(H_IMG, W_IMG))
        frames_timed = T * stage_passes  # frames behind the per-stage event times
        popc = None
        import ctypes as C
        v = C.c_double()
        if orbx.lib.orbx_popc_microbench(local_rank, 0, C.byref(v)) == 0:
            popc = v.value
        K_mean = float(n_dev.mean())
        pairs_eval = int(np.sum(n_dev[:-1] * n_dev[1:]))  # (query, train) pairs whose distance is evaluated per pass
        alg = {"pyramid": P_px, "fast": P_px + 8 * cand0, "brief": 2 * P_px + 572 * K_mean}
        stages = {}
        for name, (ms, ln) in stage.items():
            stages[name] = {"ms_per_pass": ms / stage_passes, "launches_per_pass": ln / stage_passes}
        stages["hamming"] = {"ms_per_pass": match_ms / stage_passes, "launches_per_pass": match_launches / stage_passes}
        for name, nbytes in (("pyramid", alg["pyramid"]), ("fast", alg["fast"])):
            gbs = nbytes * frames_timed / (stage[name][0] * 1e-3) / 1e9
            stages[name].update({"bound": "hbm", "achieved_gbs": gbs, "frac": gbs / hbm_peak})
        bb_ms = stage["blur"][0] + stage["brief"][0]
        gbs = alg["brief"] * frames_timed / (bb_ms * 1e-3) / 1e9
        stages["blur+brief"] = {"ms_per_pass": bb_ms / stage_passes, "bound": "hbm", "achieved_gbs": gbs,
                                "frac": gbs / hbm_peak}
        # S10.  Every (query, train) distance is evaluated ONCE per pass (the reverse search of the cross-check reads the
        # same evaluation), so pairs/s counts evaluated pairs - never forward + reverse.
        pairs_s = pairs_eval * stage_passes / (match_ms * 1e-3)
        use_mma = args.matcher != "popc"
        ham = stages["hamming"]
        ham.update({"pairs_per_s": pairs_s, "evaluated_pairs_per_pass": pairs_eval})
        if use_mma:
            # tcgen05 kind::i8: 2 * 288 int8 ops per pair (256 data + 32 key-carrying K); int8 dense peak = 2 x the bf16
            # rate on sm_100 (no int8 figure in MEASURED_PEAKS.json: 2 x the measured sustained bf16 number is used)
            tops = pairs_s * 2 * 288 / 1e12
            peak = 2.0 * peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops", 1590.0))
            ham.update({"kernel": "hamming_mma_kernel", "bound": "tensor", "achieved_tops": tops, "peak_tops": peak,
                        "frac": tops / peak, "peak_source": "2 x MEASURED_PEAKS.json bf16_tflops_sustained (%s)" % peak_kind,
                        "note": "stage time includes the int8 expansion and the merge / filter kernels; the kernel itself is bound "
                                "by its epilogue (best-2 + reverse best-1 out of TMEM): ALU pipe 75 %, tensor pipe 41 % "
                                "(profiles/r02_ncu_full_v3.csv, profiles/r02_matcher_experiments.md)",
                        "popc_equivalent": {"what": "the same pairs on the XOR+POPC kernel would need 5 POPC each",
                                            "tpopc_per_s": pairs_s * 5 / 1e12, "popc_peak_tpopc": popc / 1e12 if popc else None}})
        else:
            ham.update({"kernel": "hamming_knn_kernel", "bound": "popc", "achieved_tpopc": pairs_s * 5 / 1e12,
                        "peak_tpopc": popc / 1e12 if popc else None, "frac": (pairs_s * 5 / popc) if popc else None,
                        "executed_popc_per_pair": 5, "algorithmic_popc_per_pair": 8,
                        "algorithmic_frac": (pairs_s * 8 / popc) if popc else None})
        traffic = {}
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        except Exception:
            pass
        # dominant kernel of the step
        dom = max(("pyramid", "fast", "harris", "select", "orient", "blur", "brief", "hamming"),
                  key=lambda k: stages[k]["ms_per_pass"])
        kname = {"pyramid": "resize_exact_kernel", "fast": "fast_nms_kernel", "harris": "harris_kernel",
                 "select": "select_kernel", "orient": "orient_kernel", "blur": "blur7_kernel",
                 "brief": "brief_kernel", "hamming": ham.get("kernel", "hamming_knn_kernel")}[dom]
        tr = traffic.get(kname)
        if dom == "hamming" and use_mma:
            roof = {"kernel": kname, "bound": "tensor", "achieved": ham["achieved_tops"], "peak": ham["peak_tops"],
                    "unit": "TOP/s (int8)", "frac": ham["frac"], "traffic": tr, "peak_source": ham["peak_source"],
                    "avg_launch_ms": match_ms / max(match_launches, 1)}
        elif dom == "hamming":
            roof = {"kernel": kname, "bound": "popc", "achieved": ham["achieved_tpopc"], "peak": ham["peak_tpopc"],
                    "unit": "TPOPC/s", "frac": ham["frac"], "traffic": tr,
                    "peak_source": "orbx_popc_microbench (measured in this run; POPC is 16 lanes/clk/SM)",
                    "avg_launch_ms": match_ms / max(match_launches, 1),
                    "note": "integer XOR+POPC work: neither HBM nor tensor bound; achieved = POPC instructions issued "
                            "(5 per evaluated pair) / s, frac = POPC-pipe utilisation"}
        else:
            key = {"blur": "blur+brief", "brief": "blur+brief"}.get(dom, dom)
            sg = stages.get(key, {})
            nl = max(stage[dom][1], 1)
            roof = {"kernel": kname, "bound": "hbm", "achieved": sg.get("achieved_gbs"), "peak": hbm_peak, "unit": "GB/s",
                    "frac": sg.get("frac"), "traffic": tr, "peak_source": "MEASURED_PEAKS.json hbm_gbs (%s)" % peak_kind,
                    "avg_launch_ms": stage[dom][0] / nl,
                    "algorithmic_bytes_per_launch": (alg.get(dom, 0) * frames_timed / nl) if dom in alg else None}
            if dom == "fast":
                sm_clk = (clocks.get("sm_mhz") or 1965.0) * 1e6
                tiles = fast_tiles(H_IMG, W_IMG)
                ops = tiles * 64 * 32 * 40.0 * frames_timed / (stage["fast"][0] * 1e-3)
                roof["alu"] = {"achieved": ops / 1e12, "peak": 148 * 64 * sm_clk / 1e12, "unit": "T lane-ops/s",
                               "frac": ops / (148 * 64 * sm_clk), "what": "VIMNMX3.S16x2 of the two 9-of-16 arc networks"}
                roof["note"] = "instruction-issue / ALU-pipe bound (DRAM ~4 %): see profiles/README.md"
        cfg = make_config(args)
        h2d_step = T * H_IMG * W_IMG * P
        d2h_step = T * (4 + e2e_cap * (28 + 32 + 16) + 4) * P
        line = {
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": cfg,
            "workload_stats": {"chunk": args.chunk, "host_chunk": args.host_chunk, "pose_stage": bool(args.pose),
                               "matcher_kernel": ham.get("kernel"), "keypoints_per_frame": K_mean,
                               "fast_candidates_frame0": cand0, "good_matches_per_pair": n_good / max(T - 1, 1),
                               "good_matches_per_pair_device_pass": n_good_dev / max(T - 1, 1),
                               "ms_per_pass": ms_total / passes_all, "timed_region_s": ms_total * 1e-3},
            "hamming_pairs_per_s": pairs_s * world, "hamming_pairs_per_step": pairs_eval * P,
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d_step, "d2h_bytes_per_step": d2h_step,
                    "ms_per_step": ms_e2e / args.steps, "ms_per_pass": ms_e2e / passes_all,
                    "api": "orbx_sequence_run_host",
                    "timer": "CUDA events inside liborbx (orbx_sequence_elapsed: first upload -> last download), max over ranks",
                    "host_wall_ms_per_pass_rank0": e2e_wall_ms / passes_all,
                    "h2d_gbs_effective_per_rank": h2d_step / P / (ms_e2e / passes_all * 1e-3) / 1e9,
                    "d2h_gbs_effective_per_rank": d2h_step / P / (ms_e2e / passes_all * 1e-3) / 1e9,
                    "host_link_alone_gbs": {"h2d_min_over_ranks": h2d_min, "d2h_min_over_ranks": d2h_min},
                    **({"torch_scheduled_ms_per_pass": torch_e2e} if torch_e2e else {})},
            "gpu_launches": int(launches), "roofline": roof, "stages": stages, "clocks": clocks,
            **subs,
        }
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_baseline(frames[:args.cpu_frames], os.cpu_count() or 1)
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# ---------------------------------------------------------------------------------------------------
# --workload map: BASELINE.json configs[3] — 1 M map descriptors vs 5000 frame descriptors, query-sharded
# ---------------------------------------------------------------------------------------------------
MAP_ROWS, MAP_Q = 1_000_000, 5000


def run_map(args, rank, world, local_rank):
    """One step = one frame's map-to-frame search: the 5000 query rows are broadcast from rank 0 (NCCL), every
    rank matches its slice against the resident 1 M-row map with liborbx (k = 2), the DMatch records are
    all-gathered.  Strong scaling: the work per step is fixed (5e9 Hamming pairs)."""
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — liborbx has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import slam_rs_b200 as orbx
    from slam_rs_b200.mapmatch import MapToFrameMatcher
    synth = load_synth()
    mm = MapToFrameMatcher(device=local_rank, rank=rank, world=world, dist=dist if world > 1 else None)
    nframes = 4
    if rank == 0:
        m, _ = synth.map_and_frame_descriptors(MAP_ROWS, MAP_Q, f=0)
        frames = [synth.map_and_frame_descriptors(MAP_ROWS, MAP_Q, f=f)[1] for f in range(nframes)] \
            if args.map_fresh_frames else None
        if frames is None:
            rng = np.random.default_rng(8)
            frames = []
            for f in range(nframes):
                pick = rng.choice(MAP_ROWS, MAP_Q // 2, replace=False)
                flips = rng.random((MAP_Q // 2, 256)) < 0.1
                noisy = m[pick] ^ np.packbits(flips, axis=1, bitorder="little")
                frames.append(np.concatenate([noisy, rng.integers(0, 256, (MAP_Q - MAP_Q // 2, 32), dtype=np.uint8)]))
        mm.set_train(m)
        q_dev = [torch.from_numpy(f).to(dev) for f in frames]
        q_pin = [torch.from_numpy(f).pin_memory() for f in frames]
    else:
        mm.set_train(None, rows=MAP_ROWS)
        q_dev = [None] * nframes
        q_pin = None

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for i in range(args.warmup):
        mm.knn2(q_dev[i % nframes], nq=MAP_Q)
    barrier()
    mm.matcher.set_timing(True)
    mm.matcher.timing(reset=True)
    launches0 = orbx.launch_count()
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for i in range(args.steps):
        rec, cnt = mm.knn2(q_dev[i % nframes], nq=MAP_Q)
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    launches = orbx.launch_count() - launches0
    kern_ms, kern_launches = mm.matcher.timing(reset=True)
    mm.matcher.set_timing(False)
    # end to end: query descriptors start in pinned host memory on rank 0, the records end in host memory
    host_rec = torch.zeros((MAP_Q, 2, 16), dtype=torch.uint8).pin_memory()
    stage_q = torch.zeros((MAP_Q, 32), dtype=torch.uint8, device=dev)
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    for i in range(args.steps):
        if rank == 0:
            stage_q.copy_(q_pin[i % nframes], non_blocking=True)
        rec, cnt = mm.knn2(stage_q, nq=MAP_Q)
        if rank == 0:
            host_rec.copy_(rec, non_blocking=True)
    e3.record()
    barrier()
    ms_e2e = e2.elapsed_time(e3)
    clocks = sampler.stop()
    t = torch.tensor([ms_total, ms_e2e, kern_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e, kern_ms = float(t[0]), float(t[1]), float(t[2])
    pairs_step = float(MAP_ROWS) * MAP_Q
    if rank == 0:
        import ctypes as C
        v = C.c_double()
        popc = None
        if orbx.lib.orbx_popc_microbench(local_rank, 0, C.byref(v)) == 0:
            popc = v.value
        value = pairs_step * args.steps / (ms_total * 1e-3)
        per_rank_pairs_s = (pairs_step / world) * args.steps / (kern_ms * 1e-3)
        line = {
            "metric": "Hamming pairs/s, map-to-frame 5000 x 1M (k=2)", "value": value, "unit": "pairs/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_total / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "C4 map-to-frame: 1M map descriptors (train, resident, NCCL-broadcast once) vs "
                                   "5000 frame descriptors (query, sharded over ranks), BFMatcher Hamming k=2",
                       "map_rows": MAP_ROWS, "query_rows": MAP_Q,
                       "l2": "train set 32 MB is L2-resident by design; work is POPC/ALU bound"},
            "e2e": {"value": pairs_step * args.steps / (ms_e2e * 1e-3), "unit": "pairs/s",
                    "h2d_bytes_per_step": MAP_Q * 32, "d2h_bytes_per_step": MAP_Q * 32, "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": int(launches),
            "roofline": {"kernel": "hamming_knn_kernel", "bound": "popc", "achieved": per_rank_pairs_s * 8 / 1e12,
                         "peak": popc / 1e12 if popc else None, "unit": "TPOPC/s",
                         "frac": per_rank_pairs_s * 8 / popc if popc else None, "traffic": None,
                         "pipe_util": per_rank_pairs_s * 5 / popc if popc else None,
                         "avg_launch_ms": kern_ms / max(kern_launches, 1),
                         "note": "per-rank kernel time; 8 algorithmic POPC per pair, 5 issued (carry-save)"},
            "clocks": clocks,
        }
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def bind_near_gpu(index):
    """Pin this rank to the CPUs NVML reports as local to its GPU, so that pinned host buffers are first
    touched on that NUMA node (the end-to-end pass is host-memory / PCIe bound at 8 ranks)."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, ((os.cpu_count() or 64) + 63) // 64)
        cpus = {i * 64 + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
        return sorted(cpus)
    except Exception:
        return None


_REAL_STDOUT = None


def emit(line):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    # stdout carries exactly ONE line, the JSON: this image's NCCL prints a version banner on fd 1 from C, so fd 1 is
    # pointed at stderr for the whole run and the JSON line goes to the saved descriptor (emit)
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--frames", type=int, default=T_SEQ, help="frames of the sequence (per GPU)")
    ap.add_argument("--passes", type=int, default=10,
                    help="passes over the sequence per step (a step of 10 x 1000 frames keeps the timed region of a "
                         "--steps 20 run above 2 s)")
    ap.add_argument("--matcher", default="auto", choices=["auto", "popc", "mma"],
                    help="Hamming kernel: auto = tcgen05 int8 tensor-core kernel for batches (default), popc = XOR+POPC")
    ap.add_argument("--subs", type=int, default=1,
                    help="1: also run the sub-measurements hd (configs[2]), map (configs[3], orbx_comm over all GPUs) and "
                         "latency (configs[0]) after the headline and attach them to the line")
    ap.add_argument("--hd-frames", type=int, default=48)
    ap.add_argument("--hd-passes", type=int, default=20)
    ap.add_argument("--seqs", type=int, default=8, help="sequences per GPU of the configs[4] sub-measurement")
    ap.add_argument("--map-steps", type=int, default=20)
    ap.add_argument("--latency-iters", type=int, default=200)
    ap.add_argument("--stage-timers", type=int, default=1,
                    help="1: per-stage CUDA events inside the timed region (stages run serially); 0: timed region without "
                         "them, stage breakdown from extra passes")
    ap.add_argument("--torch-e2e", type=int, default=1,
                    help="1: also time the Python / torch-stream scheduled end-to-end pass (secondary key)")
    ap.add_argument("--size", default="kitti", choices=["kitti", "hd"],
                    help="hd: 1920x1080 frames, nfeatures=5000 (BASELINE configs[2] shape; not the headline metric)")
    ap.add_argument("--chunk", type=int, default=1000, help="frames per extraction launch (device-resident pass)")
    ap.add_argument("--dual", type=int, default=1, help="1: alternate e2e pipeline stages between two streams")
    ap.add_argument("--host-chunk", type=int, default=250,
                    help="frames per pipeline stage of the end-to-end pass (H2D / kernels / D2H overlap)")
    ap.add_argument("--cpu-frames", type=int, default=500, help="frames of the sequence timed for cpu_baseline")
    ap.add_argument("--ref-frames", type=int, default=250, help="frames per step of the reference arm")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--pose", action="store_true",
                    help="also run PoseEstimator (essential-matrix RANSAC + recoverPose, SURVEY 8 f2) on every pair inside the "
                         "timed region; not part of the headline metric")
    ap.add_argument("--native", action="store_true", help="(kept for old command lines: the e2e leg is always native now)")
    ap.add_argument("--numa-bind", type=int, default=1, help="N>1: bind each rank to its GPU's local CPUs")
    ap.add_argument("--workload", default="sequence", choices=["sequence", "map"],
                    help="sequence = configs[1] (default, the metric's workload); map = configs[3] map-to-frame matching")
    ap.add_argument("--map-fresh-frames", action="store_true", help="regenerate the map per query frame (slow)")
    ap.add_argument("--overlap", action="store_true",
                    help="run matching on a second stream (about +2 %% frames/s, but per-stage event times then overlap)")
    args = ap.parse_args()
    if args.size == "hd":
        global W_IMG, H_IMG, NFEAT, WORKLOAD
        W_IMG, H_IMG, NFEAT = 1920, 1080, 5000
        WORKLOAD = WORKLOAD.replace("C2 KITTI-shaped synthetic sequence 1241x376", "C3-shaped synthetic sequence 1920x1080") \
                           .replace("nfeatures=2000", "nfeatures=5000")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    elif args.workload == "map":
        run_map(args, rank, world, local_rank)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
