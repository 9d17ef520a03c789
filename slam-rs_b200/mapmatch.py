"""Map-to-frame matching over several GPUs (caller: /root/reference/src/mapping/map.rs:72-118).

``Map::find_matches`` gathers the visible map descriptors and calls
``match_descriptors(&map_desc, descriptors)`` (map.rs:107: query = map, train = frame) followed by
``filter_good_matches(.., 2.0)`` (map.rs:108).  BASELINE.json config 3 states the transposed problem
(query = 5000 frame descriptors, train = 1 M map descriptors, k = 2).  Both are brute-force Hamming
searches whose QUERY rows are independent, so the rows are sliced across ranks:

* the train set is made resident on every rank once (NCCL broadcast from rank 0),
* every rank matches its contiguous query slice with liborbx (no CPU path),
* the 16-byte DMatch records are all-gathered in rank order (= query order).

Two hosts of the same split:

* ``ShardedMatcher`` - the C ABI's own multi-GPU entry (``orbx_comm_*``, csrc/comm.cu): ONE process drives all GPUs,
  the train set is broadcast with NCCL inside liborbx and every GPU's kernel stores its DMatch records straight into
  GPU 0's buffer through peer-mapped pointers (no gather collective).  This is what a Rust / C++ host binds.
* ``MapToFrameMatcher`` - one process per GPU over ``torch.distributed`` (plumbing only): NCCL broadcast + all-gather
  of the 16-byte records.  With world == 1 no collective is issued.
"""
import ctypes as C

import numpy as np
import torch

from ._ffi import lib, check
from .feature import FeatureMatcher, DMATCH_DTYPE, _ptr
from .sharding import query_slice, gather_rows


class ShardedMatcher:
    """``orbx_comm_*``: query-sharded BFMatcher over the GPUs ``devices`` of this box, driven from one process.
    Results are identical to ``FeatureMatcher`` on one GPU (query_idx is the global row)."""

    def __init__(self, devices):
        devices = [int(d) for d in devices]
        arr = (C.c_int * len(devices))(*devices)
        self._h = C.c_void_p()
        check(lib.orbx_comm_create(arr, len(devices), C.byref(self._h)))
        self.devices = devices
        self.rows = 0

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib.orbx_comm_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def size(self):
        return int(lib.orbx_comm_size(self._h))

    def set_train(self, desc):
        """desc: uint8 [T, 32], numpy (host) or a CUDA tensor on devices[0]; one NCCL broadcast makes it resident."""
        rows = int(desc.shape[0])
        if not torch.is_tensor(desc):
            desc = np.ascontiguousarray(desc, np.uint8)
        check(lib.orbx_comm_set_train(self._h, _ptr(desc), rows))
        check(lib.orbx_comm_sync(self._h))  # the source may be reused / freed by the caller from here on
        self.rows = rows
        return rows

    def knn_match(self, query):
        """k = 2: numpy in -> (DMatch [Q, 2], counts [Q]) numpy out."""
        q = np.ascontiguousarray(query, np.uint8)
        out = np.zeros((max(len(q), 1), 2), DMATCH_DTYPE)
        cnt = np.zeros(max(len(q), 1), np.int32)
        check(lib.orbx_comm_knn_match2(self._h, _ptr(q), len(q), _ptr(out), _ptr(cnt)))
        return out[:len(q)], cnt[:len(q)]

    def match_descriptors(self, query):
        """k = 1 (matcher.rs:24-37): one DMatch per query row; empty side -> empty."""
        q = np.ascontiguousarray(query, np.uint8)
        out = np.zeros(max(len(q), 1), DMATCH_DTYPE)
        n = C.c_int(0)
        check(lib.orbx_comm_match(self._h, _ptr(q), len(q), _ptr(out), C.byref(n)))
        return out[:n.value]

    def knn_match_device(self, query, out, cnt):
        """CUDA tensors on devices[0]: query uint8 [Q, 32], out uint8 [Q, 2, 16], cnt int32 [Q]; asynchronous."""
        check(lib.orbx_comm_knn_match2(self._h, _ptr(query), int(query.shape[0]), _ptr(out), _ptr(cnt)))

    def sync(self):
        check(lib.orbx_comm_sync(self._h))

    def set_timing(self, enable=True):
        check(lib.orbx_comm_set_timing(self._h, 1 if enable else 0))

    def timing(self, reset=True):
        """(device ms of the calls, calls, Hamming-kernel ms of rank 0) since the last reset."""
        ms, calls, k0 = C.c_float(0), C.c_int(0), C.c_float(0)
        check(lib.orbx_comm_timing(self._h, C.byref(ms), C.byref(calls), C.byref(k0), 1 if reset else 0))
        return ms.value, calls.value, k0.value


class MapToFrameMatcher:
    def __init__(self, device=0, rank=0, world=1, dist=None, matcher=None):
        """``matcher``: anything with FeatureMatcher's ``match_device`` / ``filter_good_matches`` (the world-size-2
        gloo test injects a CPU stand-in, with ``device="cpu"``, to drive this class without a GPU)."""
        self.rank, self.world, self.dist = int(rank), int(world), dist
        self.on_gpu = device != "cpu"
        if self.on_gpu:
            self.device = int(device)
            self.dev = torch.device("cuda", self.device)
            self.matcher = matcher or FeatureMatcher(device=self.device, cross_check=False)
            self.stream = torch.cuda.Stream(device=self.dev)
            self.matcher.set_stream(self.stream.cuda_stream)
        else:
            if matcher is None:
                raise ValueError("MapToFrameMatcher: liborbx has no CPU path; device='cpu' only hosts an injected test matcher")
            self.dev = torch.device("cpu")
            self.matcher = matcher
            self.stream = None
        self.train = None

    def _fork(self):
        if self.on_gpu:
            self.stream.wait_stream(torch.cuda.current_stream(self.dev))

    def _join(self):
        if self.on_gpu:
            torch.cuda.current_stream(self.dev).wait_stream(self.stream)

    # ------------------------------------------------------------------ resident train set
    def set_train(self, desc, rows=None):
        """Make the train set resident on every rank.  ``desc``: uint8 [T, 32] (numpy or CUDA tensor) on rank 0;
        other ranks pass None and ``rows``."""
        if self.rank == 0:
            t = desc if torch.is_tensor(desc) else torch.from_numpy(np.ascontiguousarray(desc))
            t = t.to(self.dev).contiguous()
            rows = int(t.shape[0])
        else:
            t = torch.empty((int(rows), 32), dtype=torch.uint8, device=self.dev)
        if self.world > 1:
            self.dist.broadcast(t, src=0)
        self.train = t
        return rows

    # ------------------------------------------------------------------ one query set against the train set
    def _gather(self, part, total):
        return gather_rows(part, total, self.world, self.dist)

    def knn2(self, query, nq=None, replicate_query=True):
        """k = 2 search of ``query`` (uint8 [Q, 32] CUDA tensor, valid on rank 0) against the resident train set.
        Returns (records uint8 [Q, 2, 16] == DMatch, counts int32 [Q]) on every rank, query_idx global."""
        if self.world > 1 and replicate_query:
            if self.rank != 0:
                query = torch.empty((int(nq), 32), dtype=torch.uint8, device=self.dev)
            self.dist.broadcast(query, src=0)
        Q = int(query.shape[0])
        lo, hi = query_slice(Q, self.rank, self.world)
        n = hi - lo
        out = torch.zeros((max(n, 1), 2, 16), dtype=torch.uint8, device=self.dev)
        cnt = torch.zeros((max(n, 1),), dtype=torch.int32, device=self.dev)
        self._fork()
        if n:
            self.matcher.match_device(query[lo:hi], n, self.train, int(self.train.shape[0]), out, cnt, k=2)
        self._join()
        out, cnt = out[:n], cnt[:n]
        if lo:
            out.view(torch.int32)[:, :, 0] += lo   # query_idx is slice-local in the kernel
        return self._gather(out, Q), self._gather(cnt, Q)

    def match_train_as_query(self, frame, ratio=2.0):
        """Reference orientation (map.rs:107-108): query = the resident (map) rows, sliced over the ranks,
        train = ``frame`` (uint8 [K, 32] CUDA tensor on rank 0); k = 1 then ``filter_good_matches(ratio)``
        on the gathered list (numpy DMatch array, rank 0; other ranks get None)."""
        K = int(frame.shape[0])
        if self.world > 1 and K:
            self.dist.broadcast(frame, src=0)
        T = int(self.train.shape[0])
        lo, hi = query_slice(T, self.rank, self.world)
        n = hi - lo
        out = torch.zeros((max(n, 1), 16), dtype=torch.uint8, device=self.dev)
        cnt = torch.zeros((1,), dtype=torch.int32, device=self.dev)
        if K == 0 or T == 0:
            # matcher.rs:29-31 (reached from map.rs:107): an empty side gives an empty match list on every rank
            return np.zeros(0, DMATCH_DTYPE) if self.rank == 0 else None
        self._fork()
        if n:
            self.matcher.match_device(self.train[lo:hi], n, frame, K, out, cnt, k=1)
        self._join()
        out = out[:n]
        if lo:
            out.view(torch.int32)[:, 0] += lo
        allm = self._gather(out, T)
        if self.rank != 0:
            return None
        m = allm.cpu().numpy().view(DMATCH_DTYPE).reshape(-1)
        return self.matcher.filter_good_matches(m, ratio)
