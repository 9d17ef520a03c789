"""Multi-GPU partitioning of the hot path: one process per GPU, torch.distributed plumbing.

Extraction shards by frame / sequence with no data-path collective.  Large map-to-frame matching
(/root/reference/src/mapping/map.rs:107) shards the QUERY rows: the train set is broadcast once, every
rank matches its contiguous query slice with liborbx, and the 16-byte records are all-gathered.
"""


def query_slice(nq, rank, world):
    """Contiguous, balanced [lo, hi) slice of the query rows for this rank."""
    base, rem = divmod(int(nq), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def frame_shard(nframes, rank, world):
    """Frame indices of this rank (round-robin; a whole sequence stays on one rank in the bench)."""
    return list(range(rank, int(nframes), int(world)))


def gather_rows(part, total_rows, world, dist):
    """All-gather per-rank row blocks of unequal length into [total_rows, ...] (rank order = row order)."""
    import torch
    if world == 1:
        return part
    sizes = [query_slice(total_rows, r, world) for r in range(world)]
    maxn = max(hi - lo for lo, hi in sizes)
    pad = torch.zeros((maxn,) + tuple(part.shape[1:]), dtype=part.dtype, device=part.device)
    pad[:part.shape[0]] = part
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad)
    return torch.cat([bufs[r][:hi - lo] for r, (lo, hi) in enumerate(sizes)], 0)
