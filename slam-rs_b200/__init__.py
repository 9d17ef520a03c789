"""slam-rs_b200 — B200-native ORB extraction + Hamming matching behind slam-rs's `feature` API.

The directory name contains a hyphen, so the importable alias is the sibling package
``slam_rs_b200`` (its __init__ points __path__ here).  Everything computes in liborbx.so
(csrc/, hand-written CUDA for sm_100a); importing fails loudly if the library is not built.
"""
from ._ffi import lib, OrbxError, LIB_PATH, SYMBOLS  # noqa: F401
from .feature import OrbDetector, FeatureMatcher, KEYPOINT_DTYPE, DMATCH_DTYPE  # noqa: F401
from .mapping import Map, CameraIntrinsics, Triangulator  # noqa: F401
from .odometry import PoseEstimator  # noqa: F401


def version():
    return lib.orbx_version().decode()


def launch_count():
    return int(lib.orbx_launch_count())
