#!/usr/bin/env python
"""bench.py — ORB extract + frame-to-frame Hamming match throughput (frames/s @1241x376).

Workload (BASELINE.json configs[1]): a KITTI-shaped synthetic sequence of T=1000
grayscale 1241x376 frames per GPU, nfeatures=2000.  One step = one pass of the hot path over the
sequence: S1-S9 on every frame and the T-1 (t-1 -> t) matching problems (BFMatcher k=2 + cross-check
mutual test + Lowe ratio 0.75).  N>1: every rank owns an independent sequence (weak scaling, no
data-path collective — SURVEY.md §8e); value = all frames / max-over-ranks time.

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
            return float(j["hbm_gbs"]), "measured"
        except Exception:
            pass
    return 6650.0, "fallback"


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
def cpu_pass(frames, threads):
    """One pass of the same hot path with OpenCV on the host: returns (seconds, frames_done)."""
    import cv2
    cv2.setNumThreads(threads)
    orb = cv2.ORB_create(nfeatures=NFEAT)
    bf = cv2.BFMatcher(cv2.NORM_HAMMING, False)
    bfx = cv2.BFMatcher(cv2.NORM_HAMMING, True)
    t0 = time.perf_counter()
    prev = None
    for img in frames:
        _, desc = orb.detectAndCompute(img, None)
        if prev is not None and desc is not None and len(prev) and len(desc):
            knn = bf.knnMatch(prev, desc, k=2)
            mutual = {m.queryIdx for m in bfx.match(prev, desc)}
            _ = [p[0] for p in knn if len(p) == 2 and p[0].queryIdx in mutual
                 and np.float32(p[0].distance) < np.float32(RATIO) * np.float32(p[1].distance)]
        prev = desc
    return time.perf_counter() - t0, len(frames)


def cpu_baseline(frames_sample, threads):
    import cv2
    cpu_pass(frames_sample[:3], threads)  # warm-up
    dt, n = cpu_pass(frames_sample, threads)
    return {"value": n / dt, "unit": "frames/s", "cores": threads, "kind": "reference",
            "sample": "cv2 %s ORB_create(nfeatures=2000).detectAndCompute + BFMatcher knnMatch(k=2) + crossCheck match "
                      "+ Lowe 0.75 on the first %d frames of the same sequence, %d OpenCV threads, %.1f s"
                      % (cv2.__version__, n, threads, dt)}


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
        "config": {"workload": WORKLOAD % T_SEQ, "sample_frames_per_step": nsample},
        "cpu_baseline": {"value": val, "unit": "frames/s", "cores": threads, "kind": "reference",
                         "sample": "cv2 %s (the OpenCV code the reference binds) on %d frames per step, %d threads"
                                   % (cv2.__version__, nsample, threads)},
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


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — liborbx has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    import slam_rs_b200 as orbx
    from slam_rs_b200.frontend import SequenceFrontEnd
    synth = load_synth()

    T = args.frames
    if world > 1 and args.numa_bind:
        bind_near_gpu(local_rank)  # before the pinned buffers are first touched
    # every rank owns an independent sequence (canvas seed differs): weak scaling, no collective
    frames = synth.sequence(H_IMG, W_IMG, T, canvas_seed=100 + rank, noise_seed0=1000 + 100000 * rank)
    fe = SequenceFrontEnd(W_IMG, H_IMG, nfeatures=NFEAT, max_frames=T, chunk=args.chunk, device=local_rank,
                          ratio=RATIO, cross_check=True, overlap=args.overlap,
                          host_chunk=args.host_chunk, dual=args.dual,
                          pose=orbx.CameraIntrinsics.kitti() if args.pose else None)
    frames_pinned = torch.from_numpy(frames).pin_memory()
    frames_dev = torch.zeros((T, H_IMG, fe.pitch), dtype=torch.uint8, device=dev)
    frames_dev[:, :, :W_IMG].copy_(frames_pinned)
    host_out = fe.alloc_host_outputs()
    torch.cuda.synchronize()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident: warm-up, then K timed steps with per-stage event timers on
    for _ in range(args.warmup):
        fe.run_device(frames_dev)
    barrier()
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
    for _ in range(args.steps):
        fe.run_device(frames_dev)
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    launches = orbx.launch_count() - launches0
    stage = fe.det.timing(reset=True)
    match_ms, match_launches = fe.matcher.timing(reset=True)
    fe.det.set_timing(False)
    fe.matcher.set_timing(False)

    # ---- end to end through the public API with HOST buffers (pinned): H2D + kernels + D2H per step
    for _ in range(max(1, args.warmup // 2)):
        fe.run_host(frames_pinned, host_out)
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    for _ in range(args.steps):
        fe.run_host(frames_pinned, host_out)
    e3.record()
    barrier()
    ms_e2e = e2.elapsed_time(e3)
    clocks = sampler.stop()

    # ---- optional (--native): the same end-to-end pass scheduled by liborbx itself (orbx_sequence_run_host, the entry a
    # compiled-language host calls).  The library owns its streams, so this is timed on the host clock between device
    # synchronisations; reported beside, never instead of, the event-timed numbers above.
    native = None
    if args.native and not args.pose:
        import time
        from slam_rs_b200.frontend import NativeSequence
        ns = NativeSequence(W_IMG, H_IMG, nfeatures=NFEAT, max_frames=T, chunk=args.chunk, host_chunk=args.host_chunk,
                            device=local_rank, ratio=RATIO, cross_check=True)
        for _ in range(max(1, args.warmup // 2)):
            ns.run_host(frames_pinned, host_out)
            ns.sync()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            ns.run_host(frames_pinned, host_out)
        ns.sync()
        native_ms = (time.perf_counter() - t0) * 1e3
        barrier()
        native = {"value": T * world * args.steps / (native_ms * 1e-3), "unit": "frames/s",
                  "ms_per_step": native_ms / args.steps, "timer": "host clock between device synchronisations (this rank)",
                  "api": "orbx_sequence_run_host"}
        ns.close()

    # ---- work accounting
    n_dev = fe.n[:T].cpu().numpy().astype(np.int64)
    n_dev = np.minimum(n_dev, fe.cap)
    pairs_fwd = int(np.sum(n_dev[:-1] * n_dev[1:]))
    pairs_per_step = 2 * pairs_fwd  # forward k=2 + reverse k=1 (cross-check)
    K_mean = float(n_dev.mean())
    n_good = int(host_out["n_good"][1:T].numpy().astype(np.int64).sum())
    cand0 = 0
    try:
        fe.det.detect_and_compute_device(frames_dev[:1], fe.kps[:1], fe.desc[:1], fe.n[:1], width=W_IMG)
        cand0 = sum(len(fe.det.candidates(l, 0)[0]) for l in range(8))
    except Exception:
        pass

    t = torch.tensor([ms_total, ms_e2e], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e = float(t[0]), float(t[1])
    frames_all = T * world * args.steps
    value = frames_all / (ms_total * 1e-3)
    e2e_value = frames_all / (ms_e2e * 1e-3)

    if rank == 0:
        hbm_peak, peak_kind = measured_peaks()
        P = sum(pyramid_pixels(H_IMG, W_IMG))
        frames_timed = T * args.steps
        popc = None
        import ctypes as C
        v = C.c_double()
        if orbx.lib.orbx_popc_microbench(local_rank, 0, C.byref(v)) == 0:
            popc = v.value
        alg = {
            "pyramid": P, "fast": P + 8 * cand0, "brief": 2 * P + 572 * K_mean,
        }
        stages = {}
        for name, (ms, ln) in stage.items():
            stages[name] = {"ms_per_step": ms / args.steps, "launches_per_step": ln / args.steps}
        stages["hamming"] = {"ms_per_step": match_ms / args.steps, "launches_per_step": match_launches / args.steps}
        for name, nbytes in (("pyramid", alg["pyramid"]), ("fast", alg["fast"])):
            gbs = nbytes * frames_timed / (stage[name][0] * 1e-3) / 1e9
            stages[name].update({"bound": "hbm", "achieved_gbs": gbs, "frac": gbs / hbm_peak})
        bb_ms = stage["blur"][0] + stage["brief"][0]
        gbs = alg["brief"] * frames_timed / (bb_ms * 1e-3) / 1e9
        stages["blur+brief"] = {"ms_per_step": bb_ms / args.steps, "bound": "hbm", "achieved_gbs": gbs,
                                "frac": gbs / hbm_peak}
        # S10: the reference does two passes for k=2 + cross-check (forward k=2, reverse k=1): 2 x pairs_fwd
        # 256-bit comparisons of 8 POPC each are the ALGORITHMIC work.  The kernel evaluates every pair once
        # (fused reverse best-1) with 5 POPC (carry-save compression), so `frac` against the POPC-pipe peak can
        # exceed 1; `pipe_util` is the POPC instructions actually issued / peak.
        pairs_s = pairs_per_step * args.steps / (match_ms * 1e-3)
        exec_popc_s = pairs_fwd * 5.0 * args.steps / (match_ms * 1e-3)
        stages["hamming"].update({"bound": "popc", "pairs_per_s": pairs_s, "achieved_tpopc": pairs_s * 8 / 1e12,
                                  "peak_tpopc": popc / 1e12 if popc else None,
                                  "frac": (pairs_s * 8 / popc) if popc else None,
                                  "executed_popc_per_evaluated_pair": 5, "evaluated_pairs_per_step": pairs_fwd,
                                  "pipe_util": (exec_popc_s / popc) if popc else None})
        traffic = {}
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        except Exception:
            pass
        # dominant kernel of the step
        dom = max(("pyramid", "fast", "harris", "select", "orient", "blur", "brief", "hamming"),
                  key=lambda k: stages[k]["ms_per_step"])
        kname = {"pyramid": "resize_exact_kernel", "fast": "fast_nms_kernel", "harris": "harris_kernel",
                 "select": "select_kernel", "orient": "orient_kernel", "blur": "blur7_kernel",
                 "brief": "brief_kernel", "hamming": "hamming_knn_kernel"}[dom]
        tr = traffic.get(kname)
        if dom == "hamming":
            roof = {"kernel": kname, "bound": "popc", "achieved": pairs_s * 8 / 1e12,
                    "peak": popc / 1e12 if popc else None, "unit": "TPOPC/s",
                    "frac": (pairs_s * 8 / popc) if popc else None, "traffic": tr,
                    "pipe_util": stages["hamming"]["pipe_util"],
                    "peak_source": "orbx_popc_microbench (measured in this run; POPC is 16 lanes/clk/SM)",
                    "avg_launch_ms": match_ms / max(match_launches, 1),
                    "note": "integer XOR+POPC work: neither HBM nor tensor bound (north_star stage 5); achieved = "
                            "algorithmic 8 POPC x (forward + reverse) pairs, the kernel issues 5 POPC per pair once"}
            # the HBM view of the same kernel, for the record: every descriptor row is read once (32 B) and every query
            # row writes its two DMatch records, its count and one reverse-best word (44 B): far below the HBM roofline
            ham_bytes = float(np.sum(n_dev)) * (32 + 44) * args.steps
            ham_gbs = ham_bytes / (match_ms * 1e-3) / 1e9
            roof["hbm"] = {"bound": "hbm", "achieved": ham_gbs, "peak": hbm_peak, "unit": "GB/s", "frac": ham_gbs / hbm_peak,
                           "algorithmic_bytes_per_launch": ham_bytes / max(match_launches, 1), "traffic": tr,
                           "peak_source": "MEASURED_PEAKS.json hbm_gbs (%s)" % peak_kind}
        else:
            key = {"blur": "blur+brief", "brief": "blur+brief"}.get(dom, dom)
            s = stages.get(key, {})
            nl = max(stage[dom][1], 1)
            roof = {"kernel": kname, "bound": "hbm", "achieved": s.get("achieved_gbs"), "peak": hbm_peak, "unit": "GB/s",
                    "frac": s.get("frac"), "traffic": tr, "peak_source": "MEASURED_PEAKS.json hbm_gbs (%s)" % peak_kind,
                    "avg_launch_ms": stage[dom][0] / nl,
                    "algorithmic_bytes_per_launch": (alg.get(dom, 0) * frames_timed / nl) if dom in alg else None}
            if dom == "fast":
                # algorithmic ALU work of FAST: 40 packed 3-input min/max lane-ops per scored pixel, 609 tiles x 64x32
                # scored pixels per 1241x376 frame; ALU-pipe peak = 64 lanes/clk/SM (scripts/exp/pipe_exp.cu)
                sm_clk = (clocks.get("sm_mhz") or 1965.0) * 1e6
                tiles = fast_tiles(H_IMG, W_IMG)
                ops = tiles * 64 * 32 * 40.0 * frames_timed / (stage["fast"][0] * 1e-3)
                roof["alu"] = {"achieved": ops / 1e12, "peak": 148 * 64 * sm_clk / 1e12, "unit": "T lane-ops/s",
                               "frac": ops / (148 * 64 * sm_clk), "what": "VIMNMX3.S16x2 of the two 9-of-16 arc networks"}
                # FAST-9 needs 160 VIMNMX3.S16x2 per 4 pixels (two 9-of-16 sliding min/max networks): its floor on
                # this chip is the ALU pipe (64 lanes/clk/SM, scripts/exp/pipe_exp.cu), not HBM
                roof["note"] = ("ALU-pipe bound: 40 packed min/max per pixel; ncu sm__pipe_alu ~75 % busy, DRAM ~3 % "
                                "(profiles/)")
        line = {
            "metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD % T, "frames_per_gpu": T, "nfeatures": NFEAT, "chunk": args.chunk, "pose_stage": bool(args.pose), "host_chunk": args.host_chunk, "e2e_streams": 2 if args.dual else 1,
                       "l2": "inputs larger than L2 (%.0f MB of frames per pass)" % (T * H_IMG * fe.pitch / 1e6),
                       "keypoints_per_frame": K_mean, "fast_candidates_frame0": cand0,
                       "good_matches_per_pair": n_good / max(T - 1, 1)},
            "hamming_pairs_per_s": pairs_s * world, "hamming_pairs_per_step": pairs_per_step,
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": fe.h2d_bytes(T),
                    "d2h_bytes_per_step": fe.d2h_bytes(T), "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": int(launches), "roofline": roof, "stages": stages, "clocks": clocks,
            **({"e2e_native": native} if native else {}),
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
    ap.add_argument("--frames", type=int, default=T_SEQ, help="frames per GPU per step")
    ap.add_argument("--size", default="kitti", choices=["kitti", "hd"],
                    help="hd: 1920x1080 frames, nfeatures=5000 (BASELINE configs[2] shape; not the headline metric)")
    ap.add_argument("--chunk", type=int, default=1000, help="frames per extraction launch (device-resident pass)")
    ap.add_argument("--dual", type=int, default=1, help="1: alternate e2e pipeline stages between two streams")
    ap.add_argument("--host-chunk", type=int, default=167,
                    help="frames per pipeline stage of the end-to-end pass (H2D / kernels / D2H overlap)")
    ap.add_argument("--cpu-frames", type=int, default=500, help="frames of the sequence timed for cpu_baseline")
    ap.add_argument("--ref-frames", type=int, default=250, help="frames per step of the reference arm")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--pose", action="store_true",
                    help="also run PoseEstimator (essential-matrix RANSAC + recoverPose, SURVEY 8 f2) on every pair inside the "
                         "timed region; not part of the headline metric")
    ap.add_argument("--native", action="store_true",
                    help="also time the end-to-end pass through the library-scheduled pipeline (orbx_sequence_run_host)")
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
