"""Map-to-frame matching over several GPUs (caller: /root/reference/src/mapping/map.rs:72-118).

``Map::find_matches`` gathers the visible map descriptors and calls
``match_descriptors(&map_desc, descriptors)`` (map.rs:107: query = map, train = frame) followed by
``filter_good_matches(.., 2.0)`` (map.rs:108).  BASELINE.json config 3 states the transposed problem
(query = 5000 frame descriptors, train = 1 M map descriptors, k = 2).  Both are brute-force Hamming
searches whose QUERY rows are independent, so the rows are sliced across ranks:

* the train set is made resident on every rank once (NCCL broadcast from rank 0),
* every rank matches its contiguous query slice with liborbx (no CPU path),
* the 16-byte DMatch records are all-gathered in rank order (= query order).

One process per GPU; ``torch.distributed`` is plumbing only.  With world == 1 no collective is issued.
"""
import numpy as np
import torch

from .feature import FeatureMatcher, DMATCH_DTYPE
from .sharding import query_slice


class MapToFrameMatcher:
    def __init__(self, device=0, rank=0, world=1, dist=None):
        self.device, self.rank, self.world, self.dist = int(device), int(rank), int(world), dist
        self.dev = torch.device("cuda", self.device)
        self.matcher = FeatureMatcher(device=self.device, cross_check=False)
        self.stream = torch.cuda.Stream(device=self.dev)
        self.matcher.set_stream(self.stream.cuda_stream)
        self.train = None

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
        if self.world == 1:
            return part
        sizes = [query_slice(total, r, self.world) for r in range(self.world)]
        maxn = max(hi - lo for lo, hi in sizes)
        pad = torch.zeros((maxn,) + tuple(part.shape[1:]), dtype=part.dtype, device=part.device)
        pad[:part.shape[0]] = part
        bufs = [torch.empty_like(pad) for _ in range(self.world)]
        self.dist.all_gather(bufs, pad)
        return torch.cat([bufs[r][:hi - lo] for r, (lo, hi) in enumerate(sizes)], 0)

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
        cur = torch.cuda.current_stream(self.dev)
        self.stream.wait_stream(cur)
        if n:
            self.matcher.match_device(query[lo:hi], n, self.train, int(self.train.shape[0]), out, cnt, k=2)
        cur.wait_stream(self.stream)
        out, cnt = out[:n], cnt[:n]
        if lo:
            out.view(torch.int32)[:, :, 0] += lo   # query_idx is slice-local in the kernel
        return self._gather(out, Q), self._gather(cnt, Q)

    def match_train_as_query(self, frame, ratio=2.0):
        """Reference orientation (map.rs:107-108): query = the resident (map) rows, sliced over the ranks,
        train = ``frame`` (uint8 [K, 32] CUDA tensor on rank 0); k = 1 then ``filter_good_matches(ratio)``
        on the gathered list (numpy DMatch array, rank 0; other ranks get None)."""
        K = int(frame.shape[0])
        if self.world > 1:
            self.dist.broadcast(frame, src=0)
        T = int(self.train.shape[0])
        lo, hi = query_slice(T, self.rank, self.world)
        n = hi - lo
        out = torch.zeros((max(n, 1), 16), dtype=torch.uint8, device=self.dev)
        cnt = torch.zeros((1,), dtype=torch.int32, device=self.dev)
        cur = torch.cuda.current_stream(self.dev)
        self.stream.wait_stream(cur)
        if n and K:
            self.matcher.match_device(self.train[lo:hi], n, frame, K, out, cnt, k=1)
        cur.wait_stream(self.stream)
        out = out[:n]
        if lo:
            out.view(torch.int32)[:, 0] += lo
        allm = self._gather(out, T)
        if self.rank != 0:
            return None
        m = allm.cpu().numpy().view(DMATCH_DTYPE).reshape(-1)
        return self.matcher.filter_good_matches(m, ratio)
