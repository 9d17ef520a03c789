"""Host mirror of the hot part of ``mapping::Map`` (/root/reference/src/mapping/map.rs) over liborbx's
device-resident map store (csrc/mapstore.cu): same method names and argument meaning as the Rust type.

Positions, descriptors and observation counts live in HBM; ``find_matches`` culls by projection, gathers the
visible descriptors and runs the k=1 Hamming search + ``filter_good_matches(2.0)`` on the device — the reference
clones every visible descriptor and rebuilds a Mat byte by byte per frame (map.rs:86-104, 155-177).
"""
import ctypes as C

import numpy as np

from . import _ffi
from ._ffi import lib, check
from .feature import _ptr, FeatureMatcher


class CameraIntrinsics:
    """odometry::CameraIntrinsics (src/odometry/camera.rs:5-55): fx, fy, cx, cy; presets; ``to_matrix()``."""

    def __init__(self, fx, fy, cx, cy):
        self.fx, self.fy, self.cx, self.cy = float(fx), float(fy), float(cx), float(cy)

    @staticmethod
    def kitti():
        return CameraIntrinsics(718.856, 718.856, 607.1928, 185.2157)

    @staticmethod
    def webcam_vga():
        """camera.rs:29-36"""
        return CameraIntrinsics(500.0, 500.0, 320.0, 240.0)

    def to_matrix(self):
        """camera.rs:39-54: the 3x3 float64 camera matrix K."""
        import numpy as np
        return np.array([[self.fx, 0.0, self.cx], [0.0, self.fy, self.cy], [0.0, 0.0, 1.0]])


class Map:
    """``Map::new(intrinsics)`` (map.rs:21-28).  ``capacity`` bounds the number of points ever added."""

    def __init__(self, intrinsics, capacity=1 << 20, device=0):
        self._h = C.c_void_p()
        self.intrinsics = intrinsics
        self.capacity = int(capacity)
        check(lib.orbx_map_create(int(device), self.capacity, intrinsics.fx, intrinsics.fy, intrinsics.cx,
                                  intrinsics.cy, C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib.orbx_map_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def size(self):
        """map.rs:45-47"""
        return int(lib.orbx_map_size(self._h))

    def clear(self):
        """map.rs:147-150"""
        check(lib.orbx_map_clear(self._h))

    def add_points(self, positions, descriptors=None, has_descriptor=None, observations=None):
        """map.rs:31-37 — positions [n, 3] float64; descriptors [n, 32] uint8 or None (``MapPoint::new``);
        returns the ids given to the new points (sequential, like ``next_id``)."""
        xyz = np.ascontiguousarray(positions, np.float64).reshape(-1, 3)
        n = len(xyz)
        d = None if descriptors is None else np.ascontiguousarray(descriptors, np.uint8).reshape(n, 32)
        hd = None if has_descriptor is None else np.ascontiguousarray(has_descriptor, np.uint8).reshape(n)
        ob = None if observations is None else np.ascontiguousarray(observations, np.int32).reshape(n)
        first = C.c_int64(0)
        check(lib.orbx_map_add_points(self._h, _ptr(xyz), _ptr(d), _ptr(hd), _ptr(ob), n, C.byref(first)))
        return np.arange(first.value, first.value + n)

    @staticmethod
    def _pose(pose):
        r, t = pose
        return np.ascontiguousarray(r, np.float64).reshape(3, 3), np.ascontiguousarray(t, np.float64).reshape(3)

    def project_points(self, pose):
        """map.rs:50-69 for every slot: (xy [n, 2], visible [n] bool, alive [n] bool, has_desc [n] bool)."""
        r, t = self._pose(pose)
        n = self._used()
        xy = np.zeros((max(n, 1), 2), np.float64)
        vis = np.zeros(max(n, 1), np.uint8)
        check(lib.orbx_map_project(self._h, _ptr(r), _ptr(t), _ptr(xy), _ptr(vis)))
        vis = vis[:n]
        return xy[:n], (vis & 1) != 0, (vis & 2) != 0, (vis & 4) != 0

    def _used(self):
        first = C.c_int64(0)
        check(lib.orbx_map_add_points(self._h, None, None, None, None, 0, C.byref(first)))
        return int(first.value)

    def find_matches(self, keypoints, descriptors, pose, matcher):
        """map.rs:72-118 — returns [(map_point_id, keypoint_idx)] as an int array [m, 2].  ``keypoints`` is
        unused, exactly as in the reference (``_keypoints``)."""
        if not isinstance(matcher, FeatureMatcher):
            raise TypeError("matcher must be a FeatureMatcher")
        r, t = self._pose(pose)
        d = np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32) if not hasattr(descriptors, "data_ptr") \
            else descriptors
        nf = int(d.shape[0])
        cap = max(self._used(), 1)
        pairs = np.zeros((cap, 2), np.int32)
        n, nvis = C.c_int(0), C.c_int(0)
        check(lib.orbx_map_find_matches(self._h, matcher._h, _ptr(r), _ptr(t), _ptr(d), nf, _ptr(pairs), cap,
                                        C.byref(n), C.byref(nvis)))
        self.last_visible = nvis.value
        return pairs[:n.value]

    def update_observations(self, matches):
        """map.rs:121-127"""
        m = np.ascontiguousarray(matches, np.int32).reshape(-1, 2)
        check(lib.orbx_map_update_observations(self._h, _ptr(m), len(m)))

    def prune_outliers(self):
        """map.rs:130-137 — returns the number of removed points"""
        n = C.c_int(0)
        check(lib.orbx_map_prune_outliers(self._h, C.byref(n)))
        return n.value

    def stable_count(self):
        """len(Map::stable_points()) (map.rs:140-145)"""
        n = C.c_int(0)
        check(lib.orbx_map_stable_count(self._h, C.byref(n)))
        return n.value

    def observations(self):
        n = self._used()
        obs = np.zeros(max(n, 1), np.int32)
        alive = np.zeros(max(n, 1), np.uint8)
        check(lib.orbx_map_read_observations(self._h, _ptr(obs), _ptr(alive), n))
        return obs[:n], alive[:n] != 0


class Triangulator:
    """``mapping::Triangulator`` (/root/reference/src/mapping/triangulation.rs:52-198) over liborbx
    (csrc/triangulate.cu).  ``triangulate`` returns arrays instead of ``Vec<MapPoint>``: ``(ids, positions
    float64 [m, 3], descriptors uint8 [m, 32], has_descriptor bool [m])`` — id = index of the correspondence, like
    ``MapPoint::new(point_3d, i)`` (:140); observations start at 1 (``Map.add_points`` default)."""

    def __init__(self, intrinsics, max_points=1 << 16, device=0):
        self._h = C.c_void_p()
        self.intrinsics = intrinsics
        self.max_points = int(max_points)
        check(lib.orbx_triangulator_create(int(device), self.max_points, intrinsics.fx, intrinsics.fy, intrinsics.cx,
                                           intrinsics.cy, C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib.orbx_triangulator_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def _thresholds(self):
        a, b = C.c_double(0), C.c_double(0)
        check(lib.orbx_triangulator_get_thresholds(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    @property
    def min_parallax_deg(self):
        return self._thresholds()[0]

    @property
    def max_reproj_error(self):
        return self._thresholds()[1]

    def with_min_parallax(self, deg):
        """triangulation.rs:71-74 (builder: returns self)"""
        check(lib.orbx_triangulator_set_thresholds(self._h, float(deg), self.max_reproj_error))
        return self

    def with_max_reproj_error(self, error):
        """triangulation.rs:77-80"""
        check(lib.orbx_triangulator_set_thresholds(self._h, self.min_parallax_deg, float(error)))
        return self

    def triangulate(self, pose1, pose2, points1, points2, descriptors=None, return_points4d=False):
        """triangulation.rs:92-158.  Raises ``OrbxError`` ("Point arrays must have the same length") on mismatched
        inputs, returns empty arrays for empty inputs."""
        r1, t1 = Map._pose(pose1)
        r2, t2 = Map._pose(pose2)
        p1 = np.ascontiguousarray(points1, np.float32).reshape(-1, 2)
        p2 = np.ascontiguousarray(points2, np.float32).reshape(-1, 2)
        n1, n2 = len(p1), len(p2)
        d = None if descriptors is None else np.ascontiguousarray(descriptors, np.uint8).reshape(-1, 32)
        cap = max(n1, 1)
        ids = np.zeros(cap, np.int32)
        xyz = np.zeros((cap, 3), np.float64)
        od = np.zeros((cap, 32), np.uint8)
        hd = np.zeros(cap, np.uint8)
        p4 = np.zeros((cap, 4), np.float32) if return_points4d else None
        m = C.c_int(0)
        check(lib.orbx_triangulate(self._h, _ptr(r1), _ptr(t1), _ptr(r2), _ptr(t2), _ptr(p1), n1, _ptr(p2), n2,
                                   _ptr(d), 0 if d is None else len(d), _ptr(ids), _ptr(xyz), _ptr(od), _ptr(hd),
                                   _ptr(p4), cap, C.byref(m)))
        k = m.value
        out = (ids[:k], xyz[:k], od[:k], hd[:k] != 0)
        return out + (p4[:n1],) if return_points4d else out

    def set_stream(self, cuda_stream_ptr):
        check(lib.orbx_triangulator_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def triangulate_batch_device(self, R2, t2, pts1, pts2, n, npairs, max_rows, ids, xyz, n_out, R1=None, t1=None):
        """``triangulate`` for ``npairs`` problems, all CUDA tensors, asynchronous (``orbx_triangulate_batch``);
        first pose = identity unless ``R1`` / ``t1`` are given."""
        check(lib.orbx_triangulate_batch(self._h, _ptr(R1), _ptr(t1), _ptr(R2), _ptr(t2), _ptr(pts1), _ptr(pts2), _ptr(n),
                                         int(npairs), int(max_rows), _ptr(ids), _ptr(xyz), _ptr(n_out)))

    def transform_points_to_world(self, positions, pose):
        """examples/point_cloud.rs:522-544 — camera-frame points of ``triangulate`` to the world frame."""
        r, t = Map._pose(pose)
        xyz = np.array(positions, np.float64).reshape(-1, 3)
        check(lib.orbx_transform_points_to_world(self._h, _ptr(r), _ptr(t), _ptr(xyz), len(xyz)))
        return xyz
