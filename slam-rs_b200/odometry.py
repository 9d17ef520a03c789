"""Mirror of `odometry::PoseEstimator` (/root/reference/src/odometry/pose.rs) over liborbx.

``extract_matched_points`` (pose.rs:27-44) is the gather that turns the matcher's DMatch list and the two keypoint
arrays into the two point lists (``orbx_extract_matched_points``).  ``compute_essential_matrix`` (pose.rs:47-80) and
``recover_pose`` (pose.rs:83-114) wrap cv::findEssentialMat(RANSAC, 0.999, 1.0, 1000) and cv::recoverPose in the
reference; here they run as csrc/pose.cu, which reproduces OpenCV's procedure down to the RANSAC index draws, so the
hypothesis that wins is OpenCV's.  The sequence front end has the batched device forms.

The reference's ``Err(String)`` cases ("Insufficient points", "Failed to compute essential matrix", "Too few
inliers") are ``OrbxError`` with code ``E_DATA`` and the same text.
"""
import ctypes as C

import numpy as np

from ._ffi import lib, check
from .feature import FeatureMatcher, _ptr
from .mapping import CameraIntrinsics


class PoseEstimator:
    def __init__(self, intrinsics: CameraIntrinsics, device=0, matcher=None, max_points=1 << 14, max_problems=1):
        self.intrinsics = intrinsics
        self.min_matches = 8  # pose.rs:21
        self._matcher = matcher or FeatureMatcher(device=device, cross_check=False)
        self._h = C.c_void_p()
        self.max_points, self.max_problems = int(max_points), int(max_problems)
        check(lib.orbx_pose_estimator_create(int(device), self.max_points, self.max_problems, intrinsics.fx,
                                             intrinsics.fy, intrinsics.cx, intrinsics.cy, C.byref(self._h)))
        self.last_inliers = self.last_iterations = self.last_good = 0
        self.last_mask = self.last_pose_mask = None

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib.orbx_pose_estimator_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def set_stream(self, cuda_stream_ptr):
        check(lib.orbx_pose_estimator_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def extract_matched_points(self, kp1, kp2, matches):
        """pose.rs:27-44: (points1, points2) as float32 [n, 2]; raises on an out-of-range match index."""
        return self._matcher.extract_matched_points(kp1, kp2, matches)

    def compute_essential_matrix(self, points1, points2):
        """pose.rs:47-80 -> E float64 [3, 3] (unit Frobenius norm, like cv::findEssentialMat's).  The RANSAC inlier
        mask, inlier count and iteration count of the call are kept in ``last_mask / last_inliers / last_iterations``."""
        p1 = np.ascontiguousarray(points1, np.float32).reshape(-1, 2)
        p2 = np.ascontiguousarray(points2, np.float32).reshape(-1, 2)
        E = np.zeros((3, 3), np.float64)
        mask = np.zeros(max(len(p1), 1), np.uint8)
        ni, it = C.c_int(0), C.c_int(0)
        check(lib.orbx_compute_essential_matrix(self._h, _ptr(p1), len(p1), _ptr(p2), len(p2), _ptr(E), _ptr(mask),
                                                C.byref(ni), C.byref(it)))
        self.last_inliers, self.last_iterations, self.last_mask = ni.value, it.value, mask[:len(p1)] != 0
        return E

    def recover_pose(self, essential, points1, points2):
        """pose.rs:83-114 -> (R float64 [3, 3], t float64 [3]); ``last_good`` / ``last_pose_mask`` hold cv::recoverPose's
        return value and mask.  Raises ("Too few inliers") when fewer than 8 points pass the depth tests."""
        E = np.ascontiguousarray(essential, np.float64).reshape(3, 3)
        p1 = np.ascontiguousarray(points1, np.float32).reshape(-1, 2)
        p2 = np.ascontiguousarray(points2, np.float32).reshape(-1, 2)
        R, t = np.zeros((3, 3), np.float64), np.zeros(3, np.float64)
        mask = np.zeros(max(len(p1), 1), np.uint8)
        ng = C.c_int(0)
        rc = lib.orbx_recover_pose(self._h, _ptr(E), _ptr(p1), len(p1), _ptr(p2), len(p2), _ptr(R), _ptr(t), _ptr(mask),
                                   C.byref(ng))
        self.last_good, self.last_pose_mask = ng.value, mask[:len(p1)] != 0
        check(rc)
        return R, t

    def estimate_pose_batch_device(self, pts1, pts2, n, npairs, max_rows, E, R, t, n_inliers, n_good, status,
                                   ransac_mask=None, pose_mask=None):
        """Both steps for ``npairs`` frame pairs, all CUDA tensors, asynchronous on the handle's stream (see
        ``orbx_estimate_pose_batch``)."""
        check(lib.orbx_estimate_pose_batch(self._h, _ptr(pts1), _ptr(pts2), _ptr(n), int(npairs), int(max_rows), _ptr(E),
                                           _ptr(R), _ptr(t), _ptr(n_inliers), _ptr(n_good), _ptr(status),
                                           _ptr(ransac_mask), _ptr(pose_mask)))
