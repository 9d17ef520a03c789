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
    """odometry::CameraIntrinsics (src/odometry/camera.rs:14-26): fx, fy, cx, cy; ``kitti()`` preset."""

    def __init__(self, fx, fy, cx, cy):
        self.fx, self.fy, self.cx, self.cy = float(fx), float(fy), float(cx), float(cy)

    @staticmethod
    def kitti():
        return CameraIntrinsics(718.856, 718.856, 607.1928, 185.2157)


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
