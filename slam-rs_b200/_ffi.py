"""ctypes binding of liborbx.so — the C ABI declared in include/orbx.h.

There is no CPU fallback: if the shared library is missing this module raises ImportError with the
build command, and every create call fails when no CUDA device is present.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "liborbx.so")

OK, E_BADARG, E_CAPACITY, E_CUDA, E_NOMEM, E_DATA = 0, -1, -2, -3, -4, -5
STAGES = ("pyramid", "fast", "harris", "select", "orient", "blur", "brief")


class Keypoint(C.Structure):
    """== cv::KeyPoint (28 B)"""
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("size", C.c_float), ("angle", C.c_float),
                ("response", C.c_float), ("octave", C.c_int32), ("class_id", C.c_int32)]


class DMatch(C.Structure):
    """== cv::DMatch (16 B)"""
    _fields_ = [("query_idx", C.c_int32), ("train_idx", C.c_int32), ("img_idx", C.c_int32),
                ("distance", C.c_float)]


class Params(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("edge_threshold", C.c_int32), ("first_level", C.c_int32), ("wta_k", C.c_int32),
                ("score_type", C.c_int32), ("patch_size", C.c_int32), ("fast_threshold", C.c_int32)]


class OrbxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("liborbx error %d: %s" % (code, msg))
        self.code = code


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "liborbx.so not built (%s). Build it with `make -C %s` or `python -c 'import __graft_entry__ as g; "
            "g.build()'`; there is no CPU fallback." % (LIB_PATH, os.path.join(_HERE, "csrc")))
    lib = C.CDLL(LIB_PATH)
    vp, i, sz, f = C.c_void_p, C.c_int, C.c_size_t, C.c_float
    ip = C.POINTER(C.c_int)
    sig = {
        "orbx_default_params": (None, [C.POINTER(Params)]),
        "orbx_detector_create": (i, [C.POINTER(Params), i, i, i, i, C.POINTER(vp)]),
        "orbx_detector_destroy": (None, [vp]),
        "orbx_detector_max_features": (i, [vp]),
        "orbx_detect": (i, [vp, vp, i, i, sz, vp, i, vp]),
        "orbx_detect_and_compute": (i, [vp, vp, i, i, sz, vp, vp, i, vp]),
        "orbx_detect_and_compute_batch": (i, [vp, vp, i, i, sz, sz, i, vp, vp, i, vp]),
        "orbx_detect_and_compute_batch_bgr": (i, [vp, vp, i, i, sz, sz, i, vp, vp, i, vp]),
        "orbx_detector_sync": (i, [vp]),
        "orbx_detector_set_stream": (i, [vp, vp]),
        "orbx_detector_set_graph": (i, [vp, i]),
        "orbx_detector_level_info": (i, [vp, i, i, ip, ip, C.POINTER(sz), C.POINTER(sz), C.POINTER(vp)]),
        "orbx_detector_read_level": (i, [vp, i, i, i, vp, sz]),
        "orbx_detector_candidates": (i, [vp, i, i, vp, i, ip]),
        "orbx_detector_set_timing": (i, [vp, i]),
        "orbx_detector_timing": (i, [vp, C.POINTER(f), ip, i]),
        "orbx_map_create": (i, [i, i, C.c_double, C.c_double, C.c_double, C.c_double, C.POINTER(vp)]),
        "orbx_map_destroy": (None, [vp]),
        "orbx_map_size": (i, [vp]),
        "orbx_map_clear": (i, [vp]),
        "orbx_map_add_points": (i, [vp, vp, vp, vp, vp, i, C.POINTER(C.c_int64)]),
        "orbx_map_project": (i, [vp, vp, vp, vp, vp]),
        "orbx_map_find_matches": (i, [vp, vp, vp, vp, vp, i, vp, i, ip, ip]),
        "orbx_map_update_observations": (i, [vp, vp, i]),
        "orbx_map_prune_outliers": (i, [vp, ip]),
        "orbx_map_stable_count": (i, [vp, ip]),
        "orbx_map_read_observations": (i, [vp, vp, vp, i]),
        "orbx_pose_estimator_create": (i, [i, i, i, C.c_double, C.c_double, C.c_double, C.c_double, C.POINTER(vp)]),
        "orbx_pose_estimator_destroy": (None, [vp]),
        "orbx_pose_estimator_set_stream": (i, [vp, vp]),
        "orbx_pose_estimator_sync": (i, [vp]),
        "orbx_compute_essential_matrix": (i, [vp, vp, i, vp, i, vp, vp, ip, ip]),
        "orbx_recover_pose": (i, [vp, vp, vp, i, vp, i, vp, vp, vp, ip]),
        "orbx_estimate_pose_batch": (i, [vp, vp, vp, vp, i, i, vp, vp, vp, vp, vp, vp, vp, vp]),
        "orbx_triangulator_create": (i, [i, i, C.c_double, C.c_double, C.c_double, C.c_double, C.POINTER(vp)]),
        "orbx_triangulator_destroy": (None, [vp]),
        "orbx_triangulator_set_thresholds": (i, [vp, C.c_double, C.c_double]),
        "orbx_triangulator_get_thresholds": (i, [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
        "orbx_triangulate": (i, [vp, vp, vp, vp, vp, vp, i, vp, i, vp, i, vp, vp, vp, vp, vp, i, ip]),
        "orbx_transform_points_to_world": (i, [vp, vp, vp, vp, i]),
        "orbx_triangulate_batch": (i, [vp, vp, vp, vp, vp, vp, vp, vp, i, i, vp, vp, vp]),
        "orbx_triangulator_set_stream": (i, [vp, vp]),
        "orbx_matcher_create": (i, [i, i, C.POINTER(vp)]),
        "orbx_matcher_destroy": (None, [vp]),
        "orbx_matcher_sync": (i, [vp]),
        "orbx_matcher_set_stream": (i, [vp, vp]),
        "orbx_match": (i, [vp, vp, i, vp, i, vp, vp]),
        "orbx_knn_match2": (i, [vp, vp, i, vp, i, vp, vp]),
        "orbx_knn_match2_static": (i, [i, vp, i, vp, i, vp, vp, vp]),
        "orbx_knn_match2_batch": (i, [vp, vp, sz, vp, vp, sz, vp, i, i, vp, vp]),
        "orbx_match_batch": (i, [vp, vp, sz, vp, vp, sz, vp, i, i, vp, vp]),
        "orbx_ratio_test": (i, [vp, vp, vp, i, f, vp, vp]),
        "orbx_ratio_test_batch": (i, [vp, vp, vp, vp, i, i, f, vp, vp]),
        "orbx_filter_good_matches": (i, [vp, vp, i, f, vp, vp]),
        "orbx_filter_good_matches_batch": (i, [vp, vp, vp, i, i, f, vp, vp]),
        "orbx_extract_matched_points": (i, [vp, vp, i, vp, i, vp, i, vp, vp]),
        "orbx_extract_matched_points_batch": (i, [vp, vp, vp, vp, vp, C.c_size_t, vp, vp, i, i, vp, vp, vp]),
        "orbx_matcher_set_mode": (i, [vp, i]),
        "orbx_matcher_set_timing": (i, [vp, i]),
        "orbx_matcher_timing": (i, [vp, C.POINTER(f), ip, i]),
        "orbx_sequence_create": (i, [i, i, i, i, i, i, i, f, i, C.POINTER(vp)]),
        "orbx_sequence_destroy": (None, [vp]),
        "orbx_sequence_capacity": (i, [vp]),
        "orbx_sequence_run_device": (i, [vp, vp, sz, sz, i]),
        "orbx_sequence_run_host": (i, [vp, vp, i, vp, vp, vp, vp, vp]),
        "orbx_sequence_results": (i, [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]),
        "orbx_sequence_sync": (i, [vp]),
        "orbx_sequence_wait_pass": (i, [vp, i]),
        "orbx_sequence_set_reference_filter": (i, [vp, i, f]),
        "orbx_sequence_set_timing": (i, [vp, i]),
        "orbx_sequence_elapsed": (i, [vp, C.POINTER(f), ip, i]),
        "orbx_comm_create": (i, [ip, i, C.POINTER(vp)]),
        "orbx_comm_destroy": (None, [vp]),
        "orbx_comm_size": (i, [vp]),
        "orbx_comm_set_train": (i, [vp, vp, i]),
        "orbx_comm_knn_match2": (i, [vp, vp, i, vp, vp]),
        "orbx_comm_match": (i, [vp, vp, i, vp, vp]),
        "orbx_comm_sync": (i, [vp]),
        "orbx_comm_set_timing": (i, [vp, i]),
        "orbx_comm_timing": (i, [vp, C.POINTER(f), ip, C.POINTER(f), i]),
        "orbx_last_error": (C.c_char_p, []),
        "orbx_version": (C.c_char_p, []),
        "orbx_launch_count": (C.c_uint64, []),
        "orbx_popc_microbench": (i, [i, i, C.POINTER(C.c_double)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)  # AttributeError here == header/library mismatch
        fn.restype = res
        fn.argtypes = args
    return lib, tuple(sig)


lib, SYMBOLS = _load()


def check(rc, allow_capacity=False):
    if rc == OK or (allow_capacity and rc == E_CAPACITY):
        return rc
    raise OrbxError(rc, lib.orbx_last_error().decode("utf-8", "replace"))


def last_error():
    return lib.orbx_last_error().decode("utf-8", "replace")
