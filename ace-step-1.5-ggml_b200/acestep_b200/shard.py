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


def time_split_plan(n_frames: int, rank: int, world: int, halo: int = 16):
    """Time-split of ONE long clip's VAE decode: rank r decodes latent frames [w0, w1) = its core
    [c0, c1) widened by `halo` frames per side (clamped to the clip, so the clip ends keep the reference's zero padding)
    and keeps the core.  halo >= 10 covers the decoder's receptive field (conv1 +-3 frames, block-0 residual units
    +-3.9, later blocks < 1, conv-transpose taps ~1.2: about 9.5 frames per side), so the stitched result equals the
    un-split decode — the same argument that makes the reference's chunk 128 / overlap 32 tiling exact
    (acestep_ggml.cpp:2114-2229).  Returns (w0, w1, c0, c1)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    c0 = rank * n_frames // world
    c1 = (rank + 1) * n_frames // world
    return max(0, c0 - halo), min(n_frames, c1 + halo), c0, c1


def gather_time_slices(local_audio: np.ndarray, n_frames: int, hop: int, dist=None) -> Optional[np.ndarray]:
    """concatenate the per-rank core slices [ (c1-c0)*hop, channels ] in rank order on rank 0 (the one exchange step)"""
    if dist is None or dist.get_world_size() == 1:
        return local_audio
    import torch

    world, rank = dist.get_world_size(), dist.get_rank()
    lens = [((r + 1) * n_frames // world - r * n_frames // world) * hop for r in range(world)]
    mx = max(lens)
    pad = np.zeros((mx,) + tuple(local_audio.shape[1:]), dtype=local_audio.dtype)
    pad[: local_audio.shape[0]] = local_audio
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
    t = torch.from_numpy(pad).to(dev)
    bufs = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
    dist.gather(t, bufs, dst=0)
    if rank != 0:
        return None
    return np.concatenate([bufs[r].cpu().numpy()[: lens[r]] for r in range(world)], axis=0)


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


class PeerSplitDecode:
    """Time-split VAE decode of ONE clip over the ranks of a torch.distributed job (one process per GPU), with the exchange step
    fused into the decode: rank 0 owns the assembled clip on its GPU and hands its CUDA IPC handle to the other ranks once; each
    rank's decode then stores its core samples straight into that buffer over NVLink (the last conv kernel's own stores), and the
    only collective left is the barrier that tells rank 0 the clip is complete."""

    def __init__(self, bridge, n_frames: int, rank: int, world: int, dist, halo: int = 16):
        import torch
        self.br, self.n, self.rank, self.world, self.dist, self.halo = bridge, int(n_frames), rank, world, dist, halo
        dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else torch.device("cpu")
        if rank == 0:
            handle = bridge.peer_buffer_create(self.n)
            t = torch.frombuffer(bytearray(handle), dtype=torch.uint8).to(dev)
        else:
            t = torch.empty(64, dtype=torch.uint8, device=dev)
        dist.broadcast(t, src=0)
        if rank != 0:
            bridge.peer_buffer_open(bytes(t.cpu().numpy().tobytes()), self.n)
        self.plan = time_split_plan(self.n, rank, world, halo)

    def decode(self, latents: np.ndarray, fetch: bool = False) -> Optional[np.ndarray]:
        """every rank passes the same full latents [n_frames, C] (3.8 MB for 600 s: replicating it is cheaper than a halo exchange);
        returns the assembled clip on rank 0 when fetch is set"""
        w0, w1, c0, c1 = self.plan
        if c1 > c0:
            self.br.vae_decode_window(np.ascontiguousarray(latents[w0:w1]), c0 - w0, c1 - c0, c0)
        self.dist.barrier()  # every rank's stores have landed (vae_decode_window synchronises its stream before returning)
        if fetch and self.rank == 0:
            return self.br.peer_buffer_read(self.n)
        return None
