"""Mirror of the first step of `odometry::PoseEstimator` (/root/reference/src/odometry/pose.rs).

Only ``extract_matched_points`` (pose.rs:27-44) is on the data-parallel path: it is the gather that turns the
matcher's DMatch list and the two keypoint arrays into the two point lists every later stage consumes
(visual_odometry.rs:111-112).  It runs as a liborbx kernel (``orbx_extract_matched_points``); the sequence front
end has the batched device form (``SequenceFrontEnd(points=True)``).

``compute_essential_matrix`` / ``recover_pose`` (pose.rs:47-114) wrap OpenCV's RANSAC five-point solver and are
out of scope (the hypothesis search is not bit-reproducible outside OpenCV); they raise.
"""
from .feature import FeatureMatcher
from .mapping import CameraIntrinsics


class PoseEstimator:
    def __init__(self, intrinsics: CameraIntrinsics, device=0, matcher=None):
        self.intrinsics = intrinsics
        self.min_matches = 8  # pose.rs:21
        self._matcher = matcher or FeatureMatcher(device=device, cross_check=False)

    def extract_matched_points(self, kp1, kp2, matches):
        """pose.rs:27-44: (points1, points2) as float32 [n, 2]; raises on an out-of-range match index."""
        return self._matcher.extract_matched_points(kp1, kp2, matches)

    def compute_essential_matrix(self, points1, points2):
        raise NotImplementedError("pose.rs:47-80 (cv::findEssentialMat RANSAC) is outside the B200 hot path")

    def recover_pose(self, essential, points1, points2):
        raise NotImplementedError("pose.rs:83-114 (cv::recoverPose) is outside the B200 hot path")
