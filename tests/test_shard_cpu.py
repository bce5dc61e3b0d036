"""CPU: the N>1 host logic (acestep_b200/shard.py) — partition properties, and a world-size-2 gloo run that shards a batch
of songs, "processes" each rank's slice independently (no data-path collective) and gathers results on rank 0."""
import os
import socket
import sys

import numpy as np
import pytest

from acestep_b200 import shard


@pytest.mark.parametrize("n,world", [(8, 1), (8, 2), (8, 8), (5, 2), (3, 8), (0, 4)])
def test_partition_covers_every_song_once(n, world):
    seen = []
    for r in range(world):
        idx = shard.song_indices(n, r, world)
        assert all(i % world == r for i in idx)
        seen += idx
    assert sorted(seen) == list(range(n))
    # CFG doubles the evaluations of a rank but never splits a song's pair across ranks
    assert sum(shard.evaluations_per_step(n, r, world, True) for r in range(world)) == 2 * n


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 5
    songs = np.arange(n * 6, dtype=np.float32).reshape(n, 3, 2)
    mine = shard.shard(songs, rank, world)
    out = shard.gather_results(mine * 2.0 + 1.0, n, dist)  # stand-in for the per-rank sampler + decode
    if rank == 0:
        q.put(out)
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_shard_and_gather():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    ref = np.arange(5 * 6, dtype=np.float32).reshape(5, 3, 2) * 2.0 + 1.0
    assert np.array_equal(out, ref)
