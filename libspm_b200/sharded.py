"""Multi-GPU driver: one process per GPU, the text / window array / pattern set is
sharded with NO collective on the search path; the only exchange is the final gather of the small
hit lists (torch.distributed: NCCL on GPUs, gloo in the CPU tests).

Two ways to shard a plain text:
  * replicated store:  ctx.set_shard(rank, world)  -> every rank scans its super-tile range of the
    same store, hits come back in global coordinates;
  * sliced text:       slice_for_rank() gives the rank its owned end positions [own_lo, own_hi) plus
    window_size-1 symbols of left context; localize()/globalize() convert the hit coordinates.
Pattern sets (config C5) may be sharded instead with pattern_slice().
"""
from __future__ import annotations

import numpy as np

from . import HIT_DTYPE


def slice_for_rank(n: int, rank: int, world: int, window_size: int):
    """contiguous owned range of END positions (exclusive ends in (own_lo, own_hi]) for this rank and
    the text slice [start, stop) it has to hold: own range + window_size-1 symbols of left context
    (long-context row of SURVEY.md section 5)."""
    per = (n + world - 1) // world
    own_lo = min(n, rank * per)
    own_hi = min(n, own_lo + per)
    start = max(0, own_lo - max(window_size - 1, 0))
    return start, own_hi, own_lo, own_hi


def globalize(hits: np.ndarray, start: int, own_lo: int) -> np.ndarray:
    """hits of a sliced store (local coordinates) -> global coordinates, keeping only the ends this
    rank owns (a hit ending inside the left context belongs to the previous rank)."""
    h = hits.copy()
    h["end"] += np.uint64(start)
    return h[h["end"] > np.uint64(own_lo)]


def pattern_slice(n_patterns: int, rank: int, world: int):
    per = (n_patterns + world - 1) // world
    lo = min(n_patterns, rank * per)
    return lo, min(n_patterns, lo + per)


def sort_hits(h: np.ndarray) -> np.ndarray:
    return h[np.lexsort((h["end"], h["seq_id"], h["pattern_id"]))]


def gather_hits(local: np.ndarray, dst: int = 0, group=None, device=None):
    """final gather of the variable-length hit lists to rank `dst` (None on the other ranks).
    Two small collectives: all_gather of the counts, then all_gather of the padded records."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return sort_hits(local)
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    dev = device if device is not None else torch.device("cpu")
    cnt = torch.tensor([len(local)], dtype=torch.int64, device=dev)
    counts = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(counts, cnt, group=group)
    counts = [int(c.item()) for c in counts]
    mx = max(max(counts), 1)
    buf = torch.zeros(mx * HIT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    if len(local):
        raw = torch.from_numpy(np.ascontiguousarray(local).view(np.uint8).copy())
        buf[:raw.numel()] = raw.to(dev)
    bufs = [torch.zeros_like(buf) for _ in range(world)]
    dist.all_gather(bufs, buf, group=group)
    if rank != dst:
        return None
    parts = [b.cpu().numpy()[:c * HIT_DTYPE.itemsize].view(HIT_DTYPE) for b, c in zip(bufs, counts)]
    return sort_hits(np.concatenate(parts))
