"""ctypes loader of the plain-C S10 restatement (oracle/hamming_oracle.c).  Test infrastructure only."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libhamming_oracle.so")


def _lib():
    if not os.path.exists(_SO):
        subprocess.check_call(["make", "-C", _HERE], stdout=subprocess.DEVNULL)
    lib = C.CDLL(_SO)
    lib.hamming_knn.restype = None
    lib.hamming_knn.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    return lib


def knn(q, t, k):
    q = np.ascontiguousarray(q, np.uint8)
    t = np.ascontiguousarray(t, np.uint8)
    idx = np.full((len(q), k), -1, np.int32)
    dist = np.full((len(q), k), -1, np.int32)
    if len(q):
        _lib().hamming_knn(q.ctypes.data, len(q), t.ctypes.data if len(t) else None, len(t), k, idx.ctypes.data,
                           dist.ctypes.data)
    return idx, dist, np.full(len(q), min(k, len(t)), np.int32)


def knn_mt(q, t, k, threads=None):
    """knn() over query chunks on several host threads (ctypes drops the GIL): the full-size checks of tests/."""
    from concurrent.futures import ThreadPoolExecutor
    threads = threads or min(32, os.cpu_count() or 1)
    q = np.ascontiguousarray(q, np.uint8)
    if len(q) < 4 * threads:
        return knn(q, t, k)
    bounds = np.linspace(0, len(q), threads + 1).astype(int)
    with ThreadPoolExecutor(threads) as ex:
        parts = list(ex.map(lambda ab: knn(q[ab[0]:ab[1]], t, k), zip(bounds[:-1], bounds[1:])))
    return (np.concatenate([p[0] for p in parts]), np.concatenate([p[1] for p in parts]),
            np.concatenate([p[2] for p in parts]))
