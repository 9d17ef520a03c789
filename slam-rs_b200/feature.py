"""Host-side mirror of the reference's ``feature`` module over the liborbx C ABI.

Same names, argument meaning and error behaviour as /root/reference/src/feature/detector.rs:15-44
(``OrbDetector``) and src/feature/matcher.rs:15-66 (``FeatureMatcher``).  Result types are numpy
structured arrays with the layout of cv::KeyPoint / cv::DMatch and a K x 32 uint8 descriptor matrix
(== cv::Mat CV_8U), which is what odometry/mapping consume (pose.rs:36-41, triangulation.rs:144-152).

The reference's Rust toolchain is absent from this image; the Rust FFI crate that binds the same
C ABI is shipped as source under rust/ (see INTEGRATION.md).  This Python mirror is what the parity
tests and the benchmark drive.  Everything here calls CUDA through liborbx — there is no CPU path.
"""
import ctypes as C

import numpy as np

from . import _ffi
from ._ffi import lib, check

KEYPOINT_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                           ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
DMATCH_DTYPE = np.dtype([("query_idx", "<i4"), ("train_idx", "<i4"), ("img_idx", "<i4"), ("distance", "<f4")])
assert KEYPOINT_DTYPE.itemsize == 28 and DMATCH_DTYPE.itemsize == 16


def _ptr(a):
    """void* of a numpy array (host) or of anything exposing data_ptr() (torch CUDA tensor)."""
    if a is None:
        return None
    if hasattr(a, "data_ptr"):
        return C.c_void_p(a.data_ptr())
    return C.c_void_p(a.ctypes.data)


def _as_gray(image):
    img = np.asarray(image)
    if img.ndim == 3:  # OpenCV gray-converts non-8UC1 input internally
        if img.shape[2] == 1:
            img = img[:, :, 0]
        elif img.shape[2] == 3 and img.dtype == np.uint8:
            return img  # BGR: converted on the device by the ingest kernel (orbx_detect_and_compute_batch_bgr)
        else:
            raise ValueError("OrbDetector expects an 8-bit single-channel or BGR image")
    if img.dtype != np.uint8 or img.ndim != 2:
        raise ValueError("OrbDetector expects an 8-bit single-channel or BGR image")
    return img


class OrbDetector:
    """``OrbDetector::new(max_features)`` (detector.rs:15-18).

    Unlike the reference wrapper, which forwards to ``ORB::create_def()`` and silently keeps
    nfeatures=500, ``max_features`` is honoured here (``reference_faithful=True`` restores the
    reference's behaviour: nfeatures stays 500 and ``max_features`` is only echoed).
    """

    def __init__(self, max_features=500, device=0, max_width=1920, max_height=1080, max_batch=1,
                 fast_threshold=20, reference_faithful=False):
        self._h = C.c_void_p()
        self._max_features = int(max_features)
        p = _ffi.Params()
        lib.orbx_default_params(C.byref(p))
        if not reference_faithful:
            p.nfeatures = int(max_features)
        p.fast_threshold = int(fast_threshold)
        self.nfeatures = p.nfeatures
        self.max_batch = int(max_batch)
        self.device = int(device)
        check(lib.orbx_detector_create(C.byref(p), device, max_width, max_height, max_batch, C.byref(self._h)))
        # OpenCV keeps ties, so a frame can yield more than nfeatures keypoints
        self.default_cap = int(self.nfeatures + self.nfeatures // 4 + 256)

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib.orbx_detector_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def max_features(self):
        """detector.rs:42-44"""
        return self._max_features

    # ------------------------------------------------------------------ host (numpy) interface
    def _run(self, image, with_desc, cap=None):
        img = np.ascontiguousarray(_as_gray(image))
        h, w = img.shape[:2]
        cap = int(cap or self.default_cap)
        while True:
            kps = np.zeros(cap, KEYPOINT_DTYPE)
            desc = np.zeros((cap, 32), np.uint8) if with_desc else None
            n = C.c_int(0)
            if img.ndim == 3:
                # 8UC3: cv::ORB gray-converts such input itself (detector.rs:22,32); here the ingest kernel does
                rc = lib.orbx_detect_and_compute_batch_bgr(self._h, _ptr(img), w, h, img.strides[0],
                                                           img.strides[0] * h, 1, _ptr(kps), _ptr(desc), cap,
                                                           C.byref(n))
            elif with_desc:
                rc = lib.orbx_detect_and_compute(self._h, _ptr(img), w, h, img.strides[0], _ptr(kps), _ptr(desc),
                                                 cap, C.byref(n))
            else:
                rc = lib.orbx_detect(self._h, _ptr(img), w, h, img.strides[0], _ptr(kps), cap, C.byref(n))
            if rc == _ffi.E_CAPACITY and n.value > cap:
                cap = n.value  # tie-keeping retainBest exceeded the buffer: retry with the needed size
                continue
            check(rc)
            k = n.value
            return kps[:k], (desc[:k] if with_desc else None)

    def detect(self, image):
        """detector.rs:20-24 — keypoints only (S1-S7)."""
        return self._run(image, False)[0]

    def detect_and_compute(self, image):
        """detector.rs:26-40 — (keypoints, K x 32 descriptors); empty image -> (0 keypoints, 0 x 32)."""
        return self._run(image, True)

    def detect_and_compute_batch(self, images, cap=None):
        """Batch of equally sized host frames [B, H, W] -> list of (keypoints, descriptors)."""
        imgs = np.ascontiguousarray(images)
        if imgs.dtype != np.uint8 or imgs.ndim != 3:
            raise ValueError("expected uint8 [B, H, W]")
        b, h, w = imgs.shape
        cap = int(cap or self.default_cap)
        kps = np.zeros((b, cap), KEYPOINT_DTYPE)
        desc = np.zeros((b, cap, 32), np.uint8)
        n = np.zeros(b, np.int32)
        check(lib.orbx_detect_and_compute_batch(self._h, _ptr(imgs), w, h, imgs.strides[1], imgs.strides[0], b,
                                                _ptr(kps), _ptr(desc), cap, _ptr(n)))
        return [(kps[i, :n[i]], desc[i, :n[i]]) for i in range(b)]

    def detect_and_compute_batch_bgr(self, images, cap=None):
        """Batch of equally sized host BGR frames [B, H, W, 3] -> list of (keypoints, descriptors)."""
        imgs = np.ascontiguousarray(images)
        if imgs.dtype != np.uint8 or imgs.ndim != 4 or imgs.shape[3] != 3:
            raise ValueError("expected uint8 [B, H, W, 3]")
        b, h, w, _ = imgs.shape
        cap = int(cap or self.default_cap)
        kps = np.zeros((b, cap), KEYPOINT_DTYPE)
        desc = np.zeros((b, cap, 32), np.uint8)
        n = np.zeros(b, np.int32)
        check(lib.orbx_detect_and_compute_batch_bgr(self._h, _ptr(imgs), w, h, imgs.strides[1], imgs.strides[0], b,
                                                    _ptr(kps), _ptr(desc), cap, _ptr(n)))
        return [(kps[i, :n[i]], desc[i, :n[i]]) for i in range(b)]

    # ------------------------------------------------------------------ device interface (torch tensors)
    def detect_and_compute_device_bgr(self, imgs, kps, desc, n, width=None):
        """Asynchronous batch over device-resident BGR frames: uint8 CUDA tensor [B, H, W(+pad), 3]."""
        b, h = int(imgs.shape[0]), int(imgs.shape[1])
        width = int(width or imgs.shape[2])
        check(lib.orbx_detect_and_compute_batch_bgr(self._h, _ptr(imgs), width, h, int(imgs.stride(1)),
                                                    int(imgs.stride(0)), b, _ptr(kps), _ptr(desc), int(kps.shape[1]),
                                                    _ptr(n)))

    def detect_and_compute_device(self, imgs, kps, desc, n, pitch=None, frame_stride=None, width=None):
        """Asynchronous batch over device-resident frames.  imgs: uint8 CUDA tensor [B, H, pitch];
        kps: uint8 CUDA tensor [B, cap, 28]; desc: uint8 CUDA [B, cap, 32] or None; n: int32 CUDA [B]."""
        b, h = int(imgs.shape[0]), int(imgs.shape[1])
        pitch = int(pitch or imgs.stride(1))
        width = int(width or imgs.shape[2])
        frame_stride = int(frame_stride or imgs.stride(0))
        cap = int(kps.shape[1])
        check(lib.orbx_detect_and_compute_batch(self._h, _ptr(imgs), width, h, pitch, frame_stride, b, _ptr(kps),
                                                _ptr(desc), cap, _ptr(n)))

    def set_stream(self, cuda_stream_ptr):
        check(lib.orbx_detector_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def set_graph(self, enable=True):
        """Single-frame host calls replay a CUDA graph (one launch instead of ~13 + copies); same bytes out."""
        check(lib.orbx_detector_set_graph(self._h, 1 if enable else 0))

    def sync(self):
        check(lib.orbx_detector_sync(self._h))

    def set_timing(self, enable=True):
        check(lib.orbx_detector_set_timing(self._h, 1 if enable else 0))

    def timing(self, reset=True):
        ms = (C.c_float * 8)()
        ln = (C.c_int * 8)()
        check(lib.orbx_detector_timing(self._h, ms, ln, 1 if reset else 0))
        return {s: (ms[i], ln[i]) for i, s in enumerate(_ffi.STAGES)}

    # ------------------------------------------------------------------ stage access (parity tests)
    def level(self, level, frame=0, blurred=False):
        """Copy one pyramid level of the last batch to host: uint8 [h, w]."""
        w, h = C.c_int(), C.c_int()
        check(lib.orbx_detector_level_info(self._h, level, 1 if blurred else 0, C.byref(w), C.byref(h), None, None,
                                           None))
        out = np.zeros((h.value, w.value), np.uint8)
        check(lib.orbx_detector_read_level(self._h, level, 1 if blurred else 0, frame, _ptr(out), out.strides[0]))
        return out

    def candidates(self, level, frame=0):
        """FAST+NMS+border-cull candidates of the last batch: (x, y, score) int arrays, sorted raster."""
        n = C.c_int(0)
        check(lib.orbx_detector_candidates(self._h, frame, level, None, 0, C.byref(n)))
        buf = np.zeros(max(n.value, 1), np.uint32)
        check(lib.orbx_detector_candidates(self._h, frame, level, _ptr(buf), n.value, C.byref(n)))
        buf = buf[:n.value]
        buf = buf[np.argsort(buf & 0xFFFFFF, kind="stable")]
        return (buf & 0xFFF).astype(np.int64), ((buf >> 12) & 0xFFF).astype(np.int64), (buf >> 24).astype(np.int64)


class FeatureMatcher:
    """``FeatureMatcher::new()`` (matcher.rs:15-21) = BFMatcher(NORM_HAMMING, crossCheck=false)."""

    MODES = {"auto": 0, "popc": 1, "mma": 2}

    def __init__(self, device=0, cross_check=False, mode="auto"):
        self._h = C.c_void_p()
        self.device = int(device)
        check(lib.orbx_matcher_create(device, 1 if cross_check else 0, C.byref(self._h)))
        if mode != "auto":
            self.set_mode(mode)

    def set_mode(self, mode):
        """Which kernel runs the Hamming search: "auto" (tensor cores for large problems, POPC for small ones),
        "popc" or "mma" (csrc/match_mma.cu: tcgen05 int8).  Results are bit-identical."""
        check(lib.orbx_matcher_set_mode(self._h, self.MODES[mode]))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib.orbx_matcher_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    @staticmethod
    def _desc(d):
        d = np.ascontiguousarray(d)
        if d.size == 0:
            return np.zeros((0, 32), np.uint8)
        if d.dtype != np.uint8 or d.ndim != 2 or d.shape[1] != 32:
            raise ValueError("descriptors must be uint8 [N, 32] (NORM_HAMMING asserts on anything else)")
        return d

    def match_descriptors(self, desc1, desc2):
        """matcher.rs:24-37 — one DMatch per query row (k=1); empty side -> empty result."""
        q, t = self._desc(desc1), self._desc(desc2)
        out = np.zeros(max(len(q), 1), DMATCH_DTYPE)
        n = C.c_int(0)
        check(lib.orbx_match(self._h, _ptr(q), len(q), _ptr(t), len(t), _ptr(out), C.byref(n)))
        return out[:n.value]

    def knn_match(self, desc1, desc2):
        """BFMatcher.knnMatch(k=2): (matches [Q, 2], n_per_query [Q])."""
        q, t = self._desc(desc1), self._desc(desc2)
        out = np.zeros((max(len(q), 1), 2), DMATCH_DTYPE)
        cnt = np.zeros(max(len(q), 1), np.int32)
        check(lib.orbx_knn_match2(self._h, _ptr(q), len(q), _ptr(t), len(t), _ptr(out), _ptr(cnt)))
        return out[:len(q)], cnt[:len(q)]

    def ratio_test(self, knn2, n_per_q, ratio=0.75):
        knn2 = np.ascontiguousarray(knn2)
        cnt = np.ascontiguousarray(n_per_q, np.int32)
        out = np.zeros(max(len(cnt), 1), DMATCH_DTYPE)
        n = C.c_int(0)
        check(lib.orbx_ratio_test(self._h, _ptr(knn2), _ptr(cnt), len(cnt), ratio, _ptr(out), C.byref(n)))
        return out[:n.value]

    def filter_good_matches(self, matches, ratio):
        """matcher.rs:40-66 — keep distance < max(ratio * min_dist, 30)."""
        m = np.ascontiguousarray(matches)
        out = np.zeros(max(len(m), 1), DMATCH_DTYPE)
        n = C.c_int(0)
        check(lib.orbx_filter_good_matches(self._h, _ptr(m), len(m), ratio, _ptr(out), C.byref(n)))
        return out[:n.value]

    def extract_matched_points(self, kp1, kp2, matches):
        """pose.rs:27-44 — (points1, points2): float32 [n, 2] arrays, points1[i] = kp1[matches[i].query_idx].pt,
        points2[i] = kp2[matches[i].train_idx].pt.  An out-of-range index raises (the reference returns Err)."""
        k1, k2 = np.ascontiguousarray(kp1, KEYPOINT_DTYPE), np.ascontiguousarray(kp2, KEYPOINT_DTYPE)
        m = np.ascontiguousarray(matches, DMATCH_DTYPE)
        p1 = np.zeros((len(m), 2), np.float32)
        p2 = np.zeros((len(m), 2), np.float32)
        check(lib.orbx_extract_matched_points(self._h, _ptr(k1), len(k1), _ptr(k2), len(k2), _ptr(m), len(m),
                                              _ptr(p1), _ptr(p2)))
        return p1, p2

    # ------------------------------------------------------------------ device / batched interface
    def extract_matched_points_batch_device(self, kp1, n1, kp2, n2, kp_stride, matches, n, npairs, max_rows, pts1,
                                            pts2, bad):
        check(lib.orbx_extract_matched_points_batch(self._h, _ptr(kp1), _ptr(n1), _ptr(kp2), _ptr(n2), kp_stride,
                                                    _ptr(matches), _ptr(n), npairs, max_rows, _ptr(pts1), _ptr(pts2),
                                                    _ptr(bad)))

    def knn_match_batch_device(self, q, q_stride, nq, t, t_stride, nt, npairs, max_rows, out, n_per_q):
        check(lib.orbx_knn_match2_batch(self._h, _ptr(q), q_stride, _ptr(nq), _ptr(t), t_stride, _ptr(nt), npairs,
                                        max_rows, _ptr(out), _ptr(n_per_q)))

    def match_batch_device(self, q, q_stride, nq, t, t_stride, nt, npairs, max_rows, out, n_out):
        check(lib.orbx_match_batch(self._h, _ptr(q), q_stride, _ptr(nq), _ptr(t), t_stride, _ptr(nt), npairs,
                                   max_rows, _ptr(out), _ptr(n_out)))

    def ratio_test_batch_device(self, knn2, n_per_q, nq, npairs, max_rows, ratio, out, n_out):
        check(lib.orbx_ratio_test_batch(self._h, _ptr(knn2), _ptr(n_per_q), _ptr(nq), npairs, max_rows, ratio,
                                        _ptr(out), _ptr(n_out)))

    def filter_good_matches_batch_device(self, matches, n, npairs, max_rows, ratio, out, n_out):
        check(lib.orbx_filter_good_matches_batch(self._h, _ptr(matches), _ptr(n), npairs, max_rows, ratio, _ptr(out),
                                                 _ptr(n_out)))

    def match_device(self, q, nq, t, nt, out, n_out=None, k=1):
        """Single problem on device-resident descriptors (CUDA tensors); asynchronous."""
        fn = lib.orbx_match if k == 1 else lib.orbx_knn_match2
        check(fn(self._h, _ptr(q), nq, _ptr(t), nt, _ptr(out), _ptr(n_out)))

    def set_stream(self, cuda_stream_ptr):
        check(lib.orbx_matcher_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def sync(self):
        check(lib.orbx_matcher_sync(self._h))

    def set_timing(self, enable=True):
        check(lib.orbx_matcher_set_timing(self._h, 1 if enable else 0))

    def timing(self, reset=True):
        ms, ln = C.c_float(0), C.c_int(0)
        check(lib.orbx_matcher_timing(self._h, C.byref(ms), C.byref(ln), 1 if reset else 0))
        return ms.value, ln.value
