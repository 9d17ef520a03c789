"""Sequence front end: the data-parallel form of the reference's per-frame loop.

/root/reference/examples/visual_odometry.rs:84-218 does, per frame, ``detect_and_compute(gray)``
(:102), ``match_descriptors(prev_desc, desc)`` (:106) and a match filter (:107), carrying one frame
of state (:207-209).  Over a whole sequence the extractions are independent and the T-1
frame-to-frame matching problems are independent once descriptors exist, so this class runs them as
batched launches of liborbx: frames -> [B] keypoints/descriptors -> [B-1] (t-1 -> t) match problems.

Matching follows the north_star composition: BFMatcher k=2 + Lowe ratio + cross-check (mutual
nearest), plus the reference's own k=1 + ``filter_good_matches(2.0)`` when ``reference_filter=True``.

torch is used only for device/pinned memory and streams; all computation is liborbx kernels.
"""
import numpy as np
import torch

from .feature import OrbDetector, FeatureMatcher, KEYPOINT_DTYPE, DMATCH_DTYPE


class SequenceFrontEnd:
    def __init__(self, width, height, nfeatures=2000, max_frames=1000, chunk=250, device=0, cap=None,
                 ratio=0.75, cross_check=True, reference_filter=False, overlap=False, host_chunk=None,
                 dual=False, points=False, pose=None, triangulate=False, pose_chunk=1024):
        self.w, self.h = int(width), int(height)
        self.device = int(device)
        self.dev = torch.device("cuda", self.device)
        self.max_frames = int(max_frames)
        self.chunk = int(min(chunk, max_frames))
        # pipeline granularity of run_host (H2D / compute / D2H overlap); run_device uses `chunk`
        self.host_chunk = int(min(host_chunk or chunk, self.chunk))
        self.nfeatures = int(nfeatures)
        self.cap = int(cap or (nfeatures + nfeatures // 8 + 64 + 31) // 32 * 32)
        self.ratio = float(ratio)
        self.reference_filter = bool(reference_filter)
        self.pitch = (self.w + 127) // 128 * 128
        torch.cuda.set_device(self.device)
        self.det = OrbDetector(nfeatures, device=device, max_width=self.w, max_height=self.h, max_batch=self.chunk)
        self.matcher = FeatureMatcher(device=device, cross_check=cross_check)
        T, cap = self.max_frames, self.cap
        u8, i32 = torch.uint8, torch.int32
        # device-resident results of one pass
        self.kps = torch.zeros((T, cap, 28), dtype=u8, device=self.dev)
        self.desc = torch.zeros((T, cap, 32), dtype=u8, device=self.dev)
        self.n = torch.zeros((T,), dtype=i32, device=self.dev)
        self.knn = torch.zeros((T, cap, 2, 16), dtype=u8, device=self.dev)
        self.n_per_q = torch.zeros((T, cap), dtype=i32, device=self.dev)
        self.good = torch.zeros((T, cap, 16), dtype=u8, device=self.dev)
        self.n_good = torch.zeros((T,), dtype=i32, device=self.dev)
        if self.reference_filter:
            self.m1 = torch.zeros((T, cap, 16), dtype=u8, device=self.dev)
            self.n_m1 = torch.zeros((T,), dtype=i32, device=self.dev)
            self.ref_good = torch.zeros((T, cap, 16), dtype=u8, device=self.dev)
            self.n_ref_good = torch.zeros((T,), dtype=i32, device=self.dev)
        # points=True: also emit PoseEstimator::extract_matched_points (pose.rs:27-44) of every pair's good matches
        # pose=CameraIntrinsics: also run compute_essential_matrix + recover_pose (pose.rs:47-114) on every pair's
        # matched points (visual_odometry.rs:111-117), batched on the device (implies points=True)
        self.pose_intrinsics = pose
        self.points = bool(points) or pose is not None
        if self.points:
            self.pts1 = torch.zeros((T, cap, 2), dtype=torch.float32, device=self.dev)
            self.pts2 = torch.zeros((T, cap, 2), dtype=torch.float32, device=self.dev)
            self.bad_index = torch.zeros((1,), dtype=i32, device=self.dev)
        if pose is not None:
            f64 = torch.float64
            self.pose_E = torch.zeros((T, 3, 3), dtype=f64, device=self.dev)
            self.pose_R = torch.zeros((T, 3, 3), dtype=f64, device=self.dev)
            self.pose_t = torch.zeros((T, 3), dtype=f64, device=self.dev)
            self.pose_inliers = torch.zeros((T,), dtype=i32, device=self.dev)
            self.pose_good = torch.zeros((T,), dtype=i32, device=self.dev)
            self.pose_status = torch.zeros((T,), dtype=i32, device=self.dev)
            # pairs per pose launch set: the solve kernel's time is one solve's latency whatever the pair count (until the
            # warp slots fill), so the whole chunk goes in one set (scratch: 0.8 MB per pair)
            self._pose_chunk = int(min(self.chunk, pose_chunk))
        # triangulate=True (needs pose): Triangulator::triangulate (triangulation.rs:92-158) of every pair between the
        # identity pose and the recovered (R, t): 3-D points in the camera frame of frame t-1, still on the device
        self.triangulate = bool(triangulate) and pose is not None
        if self.triangulate:
            self.tri_ids = torch.zeros((T, cap), dtype=i32, device=self.dev)
            self.tri_xyz = torch.zeros((T, cap, 3), dtype=torch.float64, device=self.dev)
            self.tri_n = torch.zeros((T,), dtype=i32, device=self.dev)
        # staging for the host path: frames land tightly packed (one contiguous DMA per stage); when the width is not a
        # multiple of 4 liborbx repacks them into its pitched level-0 plane with its own kernel (csrc/ingest.cu)
        self.frames_dev = torch.zeros((T, self.h, self.w), dtype=u8, device=self.dev)
        self.copy_stream = torch.cuda.Stream(device=self.dev)
        self.d2h_stream = torch.cuda.Stream(device=self.dev)
        # liborbx runs on one explicit (non-default) stream; callers' streams are joined with events
        self.stream = torch.cuda.Stream(device=self.dev)
        # overlap=True puts matching on its own stream (ordered by events) so the Hamming kernel of chunk i
        # can fill issue slots the extraction kernels of chunk i+1 leave idle: measured +2 % on B200, both
        # sides are issue-bound, so it is off by default and the per-stage event timers stay meaningful
        self.overlap = bool(overlap)
        self.mstream = torch.cuda.Stream(device=self.dev) if self.overlap else self.stream
        self.det.set_stream(self.stream.cuda_stream)
        self.matcher.set_stream(self.mstream.cuda_stream)
        # dual=True: run_host alternates its pipeline stages between two detector / matcher handles on two
        # streams, so the tail wave of every kernel of stage i overlaps kernels of stage i+1
        self.dual = bool(dual)
        if self.dual:
            self.det2 = OrbDetector(nfeatures, device=device, max_width=self.w, max_height=self.h,
                                    max_batch=self.host_chunk)
            self.matcher2 = FeatureMatcher(device=device, cross_check=cross_check)
            self.stream2 = torch.cuda.Stream(device=self.dev)
            self.det2.set_stream(self.stream2.cuda_stream)
            self.matcher2.set_stream(self.stream2.cuda_stream)
        if pose is not None:
            from .odometry import PoseEstimator
            # One estimator (scratch + stream) per pipeline stream.  The pose stage runs on a stream of its own, ordered
            # after the matched-points kernel by an event: its five-point solve is latency-bound (one warp per pair,
            # 2.6 % of the warp slots - profiles/r01_ncu_full_pose_v9.csv), so it overlaps the next chunk's extraction.
            self.pstream = torch.cuda.Stream(device=self.dev)
            self.pose_est = PoseEstimator(pose, device=device, matcher=self.matcher, max_points=cap,
                                          max_problems=self._pose_chunk)
            self.pose_est.set_stream(self.pstream.cuda_stream)
            if self.dual:
                self.pstream2 = torch.cuda.Stream(device=self.dev)
                self.pose_est2 = PoseEstimator(pose, device=device, matcher=self.matcher2, max_points=cap,
                                               max_problems=self._pose_chunk)
                self.pose_est2.set_stream(self.pstream2.cuda_stream)
            if self.triangulate:
                from .mapping import Triangulator
                self.tri = Triangulator(pose, max_points=self._pose_chunk * cap, device=device)
                self.tri.set_stream(self.pstream.cuda_stream)
                if self.dual:
                    self.tri2 = Triangulator(pose, max_points=self._pose_chunk * cap, device=device)
                    self.tri2.set_stream(self.pstream2.cuda_stream)

    # ------------------------------------------------------------------ stages
    def extract(self, frames_dev, t0, t1, det=None, step=None):
        """frames_dev: uint8 CUDA [T, h, pitch] (pitch >= w).  Extract frames t0..t1-1."""
        det = det or self.det
        step = step or self.chunk
        for a in range(t0, t1, step):
            b = min(a + step, t1)
            det.detect_and_compute_device(frames_dev[a:b], self.kps[a:b], self.desc[a:b], self.n[a:b], width=self.w)

    def match(self, t0, t1, matcher=None):
        """Match (t-1 -> t) for t in [max(t0,1), t1): result row t describes pair (t-1, t)."""
        a = max(t0, 1)
        if t1 <= a:
            return
        npairs = t1 - a
        cap = self.cap
        stride = cap * 32
        m = matcher or self.matcher
        m.knn_match_batch_device(self.desc[a - 1], stride, self.n[a - 1:], self.desc[a], stride, self.n[a:], npairs, cap,
                                 self.knn[a], self.n_per_q[a])
        m.ratio_test_batch_device(self.knn[a], self.n_per_q[a], self.n[a - 1:], npairs, cap, self.ratio, self.good[a],
                                  self.n_good[a:])
        if self.points:
            m.extract_matched_points_batch_device(self.kps[a - 1], self.n[a - 1:], self.kps[a], self.n[a:], cap,
                                                  self.good[a], self.n_good[a:], npairs, cap, self.pts1[a],
                                                  self.pts2[a], self.bad_index)
        if self.pose_intrinsics is not None:
            second = self.dual and m is self.matcher2
            pe = self.pose_est2 if second else self.pose_est
            ps = self.pstream2 if second else self.pstream
            ps.wait_stream(self.stream2 if second else self.mstream)   # matched points of these pairs are written
            for b0 in range(0, npairs, self._pose_chunk):
                nb = min(self._pose_chunk, npairs - b0)
                t0_ = a + b0
                pe.estimate_pose_batch_device(self.pts1[t0_], self.pts2[t0_], self.n_good[t0_:], nb, cap,
                                              self.pose_E[t0_], self.pose_R[t0_], self.pose_t[t0_],
                                              self.pose_inliers[t0_:], self.pose_good[t0_:], self.pose_status[t0_:])
                if self.triangulate:
                    tri = self.tri2 if second else self.tri
                    tri.triangulate_batch_device(self.pose_R[t0_], self.pose_t[t0_], self.pts1[t0_], self.pts2[t0_],
                                                 self.n_good[t0_:], nb, cap, self.tri_ids[t0_], self.tri_xyz[t0_],
                                                 self.tri_n[t0_:])
        if self.reference_filter:
            m.match_batch_device(self.desc[a - 1], stride, self.n[a - 1:], self.desc[a], stride, self.n[a:], npairs,
                                 cap, self.m1[a], self.n_m1[a:])
            m.filter_good_matches_batch_device(self.m1[a], self.n_m1[a:], npairs, cap, 2.0, self.ref_good[a],
                                               self.n_ref_good[a:])

    # ------------------------------------------------------------------ whole passes
    def run_device(self, frames_dev):
        """One pass over device-resident frames; results stay in HBM.  Asynchronous on the current stream."""
        T = int(frames_dev.shape[0])
        cur = torch.cuda.current_stream(self.dev)
        self.stream.wait_stream(cur)
        self.mstream.wait_stream(cur)
        for ps in self._pose_streams():
            ps.wait_stream(cur)
        for a in range(0, T, self.chunk):
            b = min(a + self.chunk, T)
            self.extract(frames_dev, a, b)
            self.mstream.wait_stream(self.stream)
            self.match(a, b)
        cur.wait_stream(self.stream)
        cur.wait_stream(self.mstream)
        for ps in self._pose_streams():
            cur.wait_stream(ps)
        return T

    def _pose_streams(self):
        if self.pose_intrinsics is None:
            return []
        return [self.pstream, self.pstream2] if self.dual else [self.pstream]

    def run_host(self, frames_pinned, out):
        """End-to-end pass over HOST frames (pinned uint8 [T, h, w]) into the pinned host buffers of
        ``out`` (see alloc_host_outputs).  H2D copies run chunk-wise on a copy stream and overlap the
        kernels of the previous chunk; each chunk's results go back on a third stream while the next chunk computes."""
        T = int(frames_pinned.shape[0])
        cur = torch.cuda.current_stream(self.dev)
        self.stream.wait_stream(cur)
        self.mstream.wait_stream(cur)
        self.copy_stream.wait_stream(cur)
        self.d2h_stream.wait_stream(cur)
        for ps in self._pose_streams():
            ps.wait_stream(cur)
        # chunk schedule: three short, doubling first chunks so that compute starts early (measured: 16.60 -> 16.43 ms
        # against two; a fourth, or a tapered end, loses again), full chunks in the middle, a short
        # last chunk so that little is left to copy back after the last kernel
        c = self.host_chunk
        bounds, a = [], 0
        for size in (max(c // 8, 1), max(c // 4, 1), max(c // 2, 1)):
            if T - a > c:
                bounds.append((a, a + size))
                a += size
        while T - a > c + c // 2:
            bounds.append((a, a + c))
            a += c
        if T - a > c:
            bounds.append((a, T - c // 2))
            a = T - c // 2
        bounds.append((a, T))
        events = []
        with torch.cuda.stream(self.copy_stream):
            for a, b in bounds:
                self.frames_dev[a:b].copy_(frames_pinned[a:b], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self.copy_stream)
                events.append(ev)
        prev_extracted = None
        for i, (a, b) in enumerate(bounds):
            if self.dual:
                st, det, m = ((self.stream, self.det, self.matcher), (self.stream2, self.det2, self.matcher2))[i & 1]
                if i < 2:
                    st.wait_stream(cur)
                st.wait_event(events[i])
                self.extract(self.frames_dev, a, b, det=det, step=self.host_chunk)
                ex = torch.cuda.Event()
                ex.record(st)
            else:
                st, m = self.mstream, self.matcher
                self.stream.wait_event(events[i])
                self.extract(self.frames_dev, a, b)
                ex = torch.cuda.Event()
                ex.record(self.stream)
            # keypoints and descriptors of this chunk go back while it is being matched ...
            self.d2h_stream.wait_event(ex)
            with torch.cuda.stream(self.d2h_stream):
                out["n"][a:b].copy_(self.n[a:b], non_blocking=True)
                out["kps"][a:b].copy_(self.kps[a:b], non_blocking=True)
                out["desc"][a:b].copy_(self.desc[a:b], non_blocking=True)
            if self.dual:
                if prev_extracted is not None:
                    st.wait_event(prev_extracted)  # descriptors of frame a-1 come from the other stream
                prev_extracted = ex
            else:
                self.mstream.wait_stream(self.stream)
            self.match(a, b, matcher=m)
            # ... and its matches while the next chunk computes
            self.d2h_stream.wait_stream(st)
            with torch.cuda.stream(self.d2h_stream):
                out["good"][a:b].copy_(self.good[a:b], non_blocking=True)
                out["n_good"][a:b].copy_(self.n_good[a:b], non_blocking=True)
            if self.pose_intrinsics is not None:
                # pose / triangulation results follow on their own stream (see __init__): copied when they are ready
                self.d2h_stream.wait_stream(self.pstream2 if (self.dual and (i & 1)) else self.pstream)
                with torch.cuda.stream(self.d2h_stream):
                    for key in self._extra_outputs():
                        out[key][a:b].copy_(getattr(self, key)[a:b], non_blocking=True)
        cur.wait_stream(self.stream)
        cur.wait_stream(self.mstream)
        if self.dual:
            cur.wait_stream(self.stream2)
        for ps in self._pose_streams():
            cur.wait_stream(ps)
        cur.wait_stream(self.d2h_stream)
        return T

    def _extra_outputs(self):
        """Names of the per-pair results of the optional stages (pose, triangulation) that run_host also returns."""
        keys = []
        if self.pose_intrinsics is not None:
            keys += ["pose_E", "pose_R", "pose_t", "pose_inliers", "pose_good", "pose_status"]
        if self.triangulate:
            keys += ["tri_ids", "tri_xyz", "tri_n"]
        return keys

    def alloc_host_outputs(self):
        T, cap = self.max_frames, self.cap
        pin = dict(pin_memory=True)
        extra = {k: torch.zeros(getattr(self, k).shape, dtype=getattr(self, k).dtype, **pin) for k in self._extra_outputs()}
        return {
            **extra,
            "n": torch.zeros((T,), dtype=torch.int32, **pin),
            "kps": torch.zeros((T, cap, 28), dtype=torch.uint8, **pin),
            "desc": torch.zeros((T, cap, 32), dtype=torch.uint8, **pin),
            "good": torch.zeros((T, cap, 16), dtype=torch.uint8, **pin),
            "n_good": torch.zeros((T,), dtype=torch.int32, **pin),
        }

    def h2d_bytes(self, T):
        return T * self.h * self.w

    def d2h_bytes(self, T):
        extra = sum(getattr(self, k)[0].numel() * getattr(self, k).element_size() for k in self._extra_outputs())
        return T * (4 + self.cap * (28 + 32 + 16) + 4 + extra)

    # ------------------------------------------------------------------ host views
    def frame_result(self, t):
        """(keypoints, descriptors) of frame t as numpy (synchronises)."""
        n = int(self.n[t].item())
        n = min(n, self.cap)
        k = self.kps[t, :n].cpu().numpy().view(KEYPOINT_DTYPE).reshape(-1)
        return k, self.desc[t, :n].cpu().numpy()

    def pair_result(self, t):
        """Good matches of pair (t-1, t) as a numpy DMatch array (synchronises)."""
        n = int(self.n_good[t].item())
        return self.good[t, :n].cpu().numpy().view(DMATCH_DTYPE).reshape(-1)

    def pair_points(self, t):
        """(points1, points2) of pair (t-1, t): float32 [n_good, 2] each (synchronises)."""
        n = int(self.n_good[t].item())
        return self.pts1[t, :n].cpu().numpy(), self.pts2[t, :n].cpu().numpy()

    def pair_pose(self, t):
        """(E, R, t, ransac_inliers, points_in_front, status) of pair (t-1, t) (synchronises); status = ORBX_POSE_*."""
        return (self.pose_E[t].cpu().numpy(), self.pose_R[t].cpu().numpy(), self.pose_t[t].cpu().numpy(),
                int(self.pose_inliers[t].item()), int(self.pose_good[t].item()), int(self.pose_status[t].item()))

    def pair_map_points(self, t):
        """(ids, positions float64 [m, 3]) triangulated from pair (t-1, t) (synchronises)."""
        m = int(self.tri_n[t].item())
        return self.tri_ids[t, :m].cpu().numpy(), self.tri_xyz[t, :m].cpu().numpy()

    def pair_knn(self, t):
        nq = min(int(self.n[t - 1].item()), self.cap)
        k = self.knn[t, :nq].cpu().numpy().view(DMATCH_DTYPE).reshape(nq, 2)
        return k, self.n_per_q[t, :nq].cpu().numpy()


class NativeSequence:
    """The same pipeline scheduled by liborbx itself (csrc/sequence.cu, ``orbx_sequence_*``): what a compiled-language
    host (include/orbx.hpp, rust/orbx) calls.  Frames and results are numpy / torch CPU arrays (pinned for overlap) or
    CUDA tensors; no torch stream is involved - the library owns its four streams."""

    def __init__(self, width, height, nfeatures=2000, max_frames=1000, chunk=1000, host_chunk=250, device=0, ratio=0.75,
                 cross_check=True):
        import ctypes as C
        from ._ffi import lib, check
        self._lib, self._check, self._C = lib, check, C
        self._h = C.c_void_p()
        self.w, self.h, self.max_frames = int(width), int(height), int(max_frames)
        check(lib.orbx_sequence_create(int(device), self.w, self.h, int(nfeatures), self.max_frames, int(chunk),
                                       int(host_chunk), float(ratio), 1 if cross_check else 0, C.byref(self._h)))
        self.cap = int(lib.orbx_sequence_capacity(self._h))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            self._lib.orbx_sequence_destroy(self._h)
            self._h = self._C.c_void_p()

    __del__ = close

    def alloc_host_outputs(self):
        T, cap = self.max_frames, self.cap
        pin = dict(pin_memory=True)
        return {"n": torch.zeros((T,), dtype=torch.int32, **pin),
                "kps": torch.zeros((T, cap, 28), dtype=torch.uint8, **pin),
                "desc": torch.zeros((T, cap, 32), dtype=torch.uint8, **pin),
                "good": torch.zeros((T, cap, 16), dtype=torch.uint8, **pin),
                "n_good": torch.zeros((T,), dtype=torch.int32, **pin)}

    def run_host(self, frames, out):
        """frames: uint8 CPU tensor [T, h, w] (contiguous; pinned for overlap); out: alloc_host_outputs().  Asynchronous."""
        assert frames.dtype == torch.uint8 and frames.is_contiguous() and tuple(frames.shape[1:]) == (self.h, self.w)
        vp = self._C.c_void_p
        self._check(self._lib.orbx_sequence_run_host(self._h, vp(frames.data_ptr()), int(frames.shape[0]),
                                                     vp(out["kps"].data_ptr()), vp(out["desc"].data_ptr()),
                                                     vp(out["n"].data_ptr()), vp(out["good"].data_ptr()),
                                                     vp(out["n_good"].data_ptr())))
        return int(frames.shape[0])

    def run_device(self, frames_dev):
        """frames_dev: uint8 CUDA tensor [T, h, pitch >= w]; results stay on the device (``results()``).  Asynchronous."""
        assert frames_dev.is_cuda and frames_dev.dtype == torch.uint8 and frames_dev.stride(2) == 1
        self._check(self._lib.orbx_sequence_run_device(self._h, self._C.c_void_p(frames_dev.data_ptr()),
                                                       int(frames_dev.stride(1)), int(frames_dev.stride(0)),
                                                       int(frames_dev.shape[0])))
        return int(frames_dev.shape[0])

    def sync(self):
        self._check(self._lib.orbx_sequence_sync(self._h))

    def wait_pass(self, back=0):
        """Block until the run call issued ``back`` calls before the most recent one has finished (downloads included);
        consecutive run_host calls are pipelined into each other, so a streaming host alternates two output sets and
        consumes sequence k with ``wait_pass(1)`` while sequence k+1 runs."""
        self._check(self._lib.orbx_sequence_wait_pass(self._h, int(back)))

    def set_reference_filter(self, enable=True, ratio=2.0):
        """The reference's own loop (visual_odometry.rs:106-107): match_descriptors (k=1) + filter_good_matches(ratio)
        instead of k=2 + cross-check + Lowe; ``good`` / ``n_good`` then hold that filter's survivors."""
        self._check(self._lib.orbx_sequence_set_reference_filter(self._h, 1 if enable else 0, float(ratio)))

    def set_timing(self, enable=True):
        self._check(self._lib.orbx_sequence_set_timing(self._h, 1 if enable else 0))

    def elapsed(self, reset=True):
        """(device milliseconds from the start of the first pass since the last reset to the end of the last, passes)."""
        ms, n = self._C.c_float(0), self._C.c_int(0)
        self._check(self._lib.orbx_sequence_elapsed(self._h, self._C.byref(ms), self._C.byref(n), 1 if reset else 0))
        return ms.value, n.value

    def results_to_host(self, T):
        """Device results of the last pass copied to numpy (synchronises): dict like alloc_host_outputs()."""
        C = self._C
        ptrs = [C.c_void_p() for _ in range(5)]
        self._check(self._lib.orbx_sequence_results(self._h, *[C.byref(p) for p in ptrs]))
        self.sync()
        from cuda.bindings import runtime as rt  # cuda-python, for a plain cudaMemcpy of raw device pointers
        cap = self.cap
        shapes = {"kps": ((T, cap, 28), np.uint8), "desc": ((T, cap, 32), np.uint8), "n": ((T,), np.int32),
                  "good": ((T, cap, 16), np.uint8), "n_good": ((T,), np.int32)}
        out = {}
        for (key, (shape, dt)), p in zip(shapes.items(), ptrs):
            a = np.zeros(shape, dt)
            err, = rt.cudaMemcpy(a.ctypes.data, p.value, a.nbytes, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost)
            if int(err) != 0:
                raise RuntimeError("cudaMemcpy failed: %s" % err)
            out[key] = a
        return out
