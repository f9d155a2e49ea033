"""CPU tests of the multi-GPU host logic: frame -> rank round-robin, ragged tails, and the
optional gather of encoded tensors to the consumer rank, run as world_size-2 (and 3) `gloo` process groups."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from mv3d_b200 import shard


def test_round_robin_is_a_partition():
    for world in (1, 2, 3, 4, 8):
        for n in (0, 1, 7, 8, 4096, 4099):
            seen = []
            for r in range(world):
                fr = list(shard.frames_of(r, world, n))
                assert all(shard.owner_of(f, world) == r for f in fr)
                assert [shard.local_index(f, world) for f in fr] == list(range(len(fr)))
                seen += fr
            assert sorted(seen) == list(range(n))
            assert shard.shard_sizes(world, n) == [len(shard.frames_of(r, world, n)) for r in range(world)]
            assert max(shard.shard_sizes(world, n), default=0) - min(shard.shard_sizes(world, n), default=0) <= 1


def test_interleave_inverts_sharding():
    world, n = 3, 11
    per_rank = [[f"frame{f}" for f in shard.frames_of(r, world, n)] for r in range(world)]
    assert shard.interleave(per_rank, n) == [f"frame{f}" for f in range(n)]


def test_batches_cover_in_order():
    fr = shard.frames_of(1, 4, 1000)
    got = [list(b) for b in shard.batches(fr, 64)]
    assert [f for b in got for f in b] == list(fr)
    assert all(len(b) == 64 for b in got[:-1]) and 0 < len(got[-1]) <= 64


def _fake_encoding(frame, shape=(4, 5, 3)):
    """Stands in for an encoded tensor: a deterministic function of the frame id (no GPU in these tests)."""
    rng = np.random.default_rng(frame)
    return torch.from_numpy(rng.random(shape, dtype=np.float32)) + frame


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, num_frames, batch, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        loaded = []

        def load(frame):
            loaded.append(frame)
            return np.full((frame % 5, 4), frame, np.float32)  # ragged, some empty

        def encode(frames, pts, offs):
            assert len(offs) == len(frames) + 1 and offs[-1] == pts.shape[0]
            for i, f in enumerate(frames):  # every sweep arrives at its own offset
                assert (pts[offs[i]:offs[i + 1]] == f).all() and offs[i + 1] - offs[i] == f % 5
            return torch.stack([_fake_encoding(f) for f in frames]), None

        stream = shard.ShardedStream(rank, world, num_frames, batch, load, encode)
        tops = [t for _, t, _ in stream]
        assert loaded == list(shard.frames_of(rank, world, num_frames))
        local = torch.cat(tops) if tops else torch.empty((0, 4, 5, 3))
        full = shard.gather_frames(local, num_frames, dst=0)
        slowest = shard.max_over_ranks(float(rank + 1))
        assert slowest == float(world)
        if rank == 0:
            assert full.shape[0] == num_frames
            for f in range(num_frames):
                assert torch.equal(full[f], _fake_encoding(f)), f
            open(os.path.join(out_dir, "ok"), "w").write("ok")
        else:
            assert full is None
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,num_frames,batch", [(2, 37, 8), (2, 64, 64), (3, 10, 4), (2, 1, 4)])
def test_sharded_stream_and_gather_gloo(tmp_path, world, num_frames, batch):
    mp.spawn(_worker, args=(world, _free_port(), num_frames, batch, str(tmp_path)), nprocs=world, join=True)
    assert (tmp_path / "ok").exists()
