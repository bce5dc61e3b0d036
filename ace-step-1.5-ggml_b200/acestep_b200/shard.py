"""Sample sharding across the GPUs of one box: the (song, CFG-branch) evaluations of a denoise step share
only read-only weights, so songs are dealt round-robin to ranks and BOTH CFG branches of a song stay on the same rank —
the CFG combine and the Euler update are local and the data path needs no collective.  Weights are replicated.

This is host logic only (no CUDA): `song_indices` is what `bench.py` and a multi-GPU caller of ``B200Bridge`` use to pick
their slice; `gather_results` is the one exchange of the job — results back to rank 0 — and runs on whatever
``torch.distributed`` backend the caller initialised (NCCL on the box, gloo in the CPU tests)."""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np


def song_indices(n_songs: int, rank: int, world: int) -> List[int]:
    """songs owned by `rank`: rank, rank + world, ...  (every song appears on exactly one rank)"""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    return list(range(rank, max(n_songs, 0), world))


def evaluations_per_step(n_songs: int, rank: int, world: int, cfg: bool) -> int:
    """network evaluations this rank runs per Euler step (B, or 2B with CFG; SURVEY 8d metric (i))"""
    return len(song_indices(n_songs, rank, world)) * (2 if cfg else 1)


def shard(arr: Optional[np.ndarray], rank: int, world: int) -> Optional[np.ndarray]:
    """slice of a [n_songs, ...] array owned by `rank` (contiguous copy; None passes through)"""
    if arr is None:
        return None
    return np.ascontiguousarray(arr[rank::world])


def gather_results(local: np.ndarray, n_songs: int, dist=None) -> Optional[np.ndarray]:
    """inverse of `shard`: rank 0 receives [n_songs, ...] in the original song order; other ranks get None.
    `dist` = an initialised torch.distributed module (None = single process)."""
    if dist is None or dist.get_world_size() == 1:
        return local
    import torch

    world, rank = dist.get_world_size(), dist.get_rank()
    counts = [len(song_indices(n_songs, r, world)) for r in range(world)]
    mx = max(counts)
    item = tuple(local.shape[1:])
    pad = np.zeros((mx,) + item, dtype=local.dtype)
    pad[: local.shape[0]] = local
    t = torch.from_numpy(pad)
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    t = t.to(dev)
    bufs = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
    dist.gather(t, bufs, dst=0)
    if rank != 0:
        return None
    out = np.empty((n_songs,) + item, dtype=local.dtype)
    for r in range(world):
        out[r::world] = bufs[r].cpu().numpy()[: counts[r]]
    return out
