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


@pytest.mark.parametrize("T,world,halo", [(15000, 8, 16), (100, 3, 16), (7, 8, 16), (64, 4, 0)])
def test_time_split_plan(T, world, halo):
    cores = []
    for r in range(world):
        w0, w1, c0, c1 = shard.time_split_plan(T, r, world, halo)
        assert 0 <= w0 <= c0 <= c1 <= w1 <= T
        assert c0 - w0 == min(halo, c0) and w1 - c1 == min(halo, T - c1)  # halo only where the clip continues
        cores.append((c0, c1))
    assert cores[0][0] == 0 and cores[-1][1] == T
    assert all(cores[i][1] == cores[i + 1][0] for i in range(world - 1))


def _worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = 5
    songs = np.arange(n * 6, dtype=np.float32).reshape(n, 3, 2)
    mine = shard.shard(songs, rank, world)
    out = shard.gather_results(mine * 2.0 + 1.0, n, dist)  # stand-in for the per-rank sampler + decode
    # time split of one clip: every rank "decodes" (here: repeats x hop) its halo'd window and keeps the core
    T, hop = 37, 3
    lat = np.arange(T, dtype=np.float32)
    w0, w1, c0, c1 = shard.time_split_plan(T, rank, world, 4)
    win = np.repeat(lat[w0:w1], hop)[:, None]
    stitched = shard.gather_time_slices(win[(c0 - w0) * hop:(c0 - w0) * hop + (c1 - c0) * hop], T, hop, dist)
    # PeerSplitDecode host logic with a stand-in bridge: the "peer buffer" is a file both ranks map, the "decode" repeats x hop
    class FakeBridge:
        hop_length, audio_channels = hop, 1

        def __init__(self, path):
            self.path, self.buf = path, None

        def peer_buffer_create(self, n_frames):
            np.zeros((n_frames * hop, 1), np.float32).tofile(self.path)
            self.buf = np.memmap(self.path, np.float32, "r+", shape=(n_frames * hop, 1))
            return os.path.basename(self.path).encode().ljust(64, b"\0")

        def peer_buffer_open(self, handle, n_frames):
            assert handle.rstrip(b"\0") == os.path.basename(self.path).encode() and len(handle) == 64
            self.buf = np.memmap(self.path, np.float32, "r+", shape=(n_frames * hop, 1))

        def vae_decode_window(self, lat_win, core0, core_frames, dst_frame):
            audio = np.repeat(lat_win[:, 0], hop)
            self.buf[dst_frame * hop:(dst_frame + core_frames) * hop, 0] = audio[core0 * hop:(core0 + core_frames) * hop]
            self.buf.flush()

        def peer_buffer_read(self, n_frames):
            return np.array(np.memmap(self.path, np.float32, "r", shape=(n_frames * hop, 1)))

    split = shard.PeerSplitDecode(FakeBridge(os.path.join(os.environ["ACE_TEST_TMP"], "peer.bin")), T, rank, world, dist, halo=4)
    peer = split.decode(lat[:, None], fetch=True)
    if rank == 0:
        q.put((out, stitched, peer))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gloo_shard_and_gather():
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    import tempfile
    os.environ["ACE_TEST_TMP"] = tempfile.mkdtemp(prefix="ace_peer_")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    out, stitched, peer = q.get(timeout=120)
    assert np.array_equal(stitched[:, 0], np.repeat(np.arange(37, dtype=np.float32), 3))
    assert np.array_equal(peer[:, 0], np.repeat(np.arange(37, dtype=np.float32), 3))  # peer-store split == un-split
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    ref = np.arange(5 * 6, dtype=np.float32).reshape(5, 3, 2) * 2.0 + 1.0
    assert np.array_equal(out, ref)
