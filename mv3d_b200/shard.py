"""Frame sharding of a sweep stream over the GPUs of one box (BASELINE.json config 4).

Each sweep is an independent unit (the reference routine is a stateless pure function), so the encode path needs
no collective: one process per GPU, frame `i` goes to rank `i mod world` -- the same round-robin the reference
gets from `multiprocessing.Pool.map` over frames (`src/data.py:495-506`, `src/preprocess_for_mv3d.py:62-72`).
The only exchange is optional and off the encode path: delivering encoded tensors to one consumer rank
(`gather_frames`, NCCL over NVLink on GPUs, gloo in the CPU tests).

torch.distributed is plumbing here; nothing in this module computes an encoding.
"""
from __future__ import annotations

from typing import Iterable, Iterator, List, Sequence, Tuple


def owner_of(frame: int, world: int) -> int:
    """Rank that encodes `frame`."""
    return int(frame) % int(world)


def frames_of(rank: int, world: int, num_frames: int) -> range:
    """Frames of a `num_frames`-long stream that `rank` encodes, in stream order."""
    return range(int(rank), int(num_frames), int(world))


def local_index(frame: int, world: int) -> int:
    """Position of `frame` among its owner's frames."""
    return int(frame) // int(world)


def shard_sizes(world: int, num_frames: int) -> List[int]:
    """Frames per rank (the first `num_frames % world` ranks hold one more)."""
    return [len(frames_of(r, world, num_frames)) for r in range(world)]


def batches(frames: Sequence[int], batch: int) -> Iterator[Sequence[int]]:
    """Consecutive groups of at most `batch` frames: one batched encode launch each."""
    for i in range(0, len(frames), batch):
        yield frames[i:i + batch]


def interleave(per_rank: Sequence[Sequence], num_frames: int) -> list:
    """Inverse of the sharding: `per_rank[r][j]` is frame `r + j*world`; returns the frames in stream order."""
    world = len(per_rank)
    out = []
    for f in range(num_frames):
        out.append(per_rank[owner_of(f, world)][local_index(f, world)])
    return out


def gather_frames(local, num_frames: int, dst: int = 0, group=None):
    """Deliver every rank's encoded frames to `dst`, in stream order, without a staging copy.

    `local`: tensor `[n_local, ...]` holding this rank's frames (`frames_of(rank, world, num_frames)`), on the
    device the process group works with (CUDA for nccl, CPU for gloo). Returns `[num_frames, ...]` on `dst`,
    None elsewhere.

    Frame f of the stream lives at `local[f // world]` on rank `f % world`, so on `dst` every frame is received
    straight into its final place `out[f]` (one point-to-point receive per frame, all posted as one batch: NCCL runs
    them as a single grouped operation over NVLink); the only local copy is dst's own share. Peak memory on `dst` is
    the result plus its own share -- not the 2x of a gather into per-rank buffers followed by an interleaving pass.
    """
    import torch.distributed as dist
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    sizes = shard_sizes(world, num_frames)
    assert local.shape[0] == sizes[rank], f"rank {rank} holds {local.shape[0]} frames, expected {sizes[rank]}"
    if num_frames == 0:
        return local.new_empty((0, *local.shape[1:])) if rank == dst else None
    local = local.contiguous()

    def peer(r):  # batch_isend_irecv takes global ranks
        return dist.get_global_rank(group, r) if group is not None else r

    ops = []
    out = None
    if rank == dst:
        out = local.new_empty((num_frames, *local.shape[1:]))
        for r in range(world):
            if r == dst:
                continue
            for j in range(sizes[r]):
                ops.append(dist.P2POp(dist.irecv, out[r + j * world], peer(r), group))
    else:
        for j in range(sizes[rank]):
            ops.append(dist.P2POp(dist.isend, local[j], peer(dst), group))
    reqs = dist.batch_isend_irecv(ops) if ops else []
    if rank == dst and sizes[dst]:
        out[dst::world] = local  # own share: one strided copy, overlaps the transfers
    for q in reqs:
        q.wait()
    return out


def max_over_ranks(value: float, device=None, group=None) -> float:
    """Slowest rank's time: multi-GPU numbers are the max over ranks of device-side timings."""
    import torch
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return float(value)
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def init_process_group_quiet(backend: str, device=None) -> None:
    """`dist.init_process_group` with stdout (the file descriptor, not just sys.stdout) pointed at stderr while the
    communicator comes up: NCCL writes a version banner to stdout from C, and bench.py's stdout must be exactly one
    JSON line."""
    import os
    import sys
    import torch.distributed as dist
    sys.stdout.flush()
    saved = os.dup(1)
    try:
        os.dup2(2, 1)
        if backend == "nccl" and device is not None:
            dist.init_process_group(backend, device_id=device)
        else:
            dist.init_process_group(backend)
        dist.barrier()  # forces communicator creation now, inside the redirected region
        if device is not None and backend == "nccl":
            import torch
            torch.cuda.synchronize(device)
    finally:
        sys.stdout.flush()
        os.dup2(saved, 1)
        os.close(saved)


class ShardedStream:
    """Drives one rank's share of a sweep stream through a `ViewEncoder`, `batch` frames per launch.

    `load(frame) -> (n, 4) float32 numpy` supplies sweeps (a `.bin` reader in production, the synthetic
    generator in bench/tests). `encode(frames, packed_points, offsets) -> (top, front)` is the encoder call;
    it is injected so the host logic can be tested without a GPU.
    """

    def __init__(self, rank: int, world: int, num_frames: int, batch: int, load, encode):
        self.rank, self.world, self.num_frames, self.batch = rank, world, num_frames, batch
        self.load, self.encode = load, encode

    def __iter__(self) -> Iterator[Tuple[Sequence[int], object, object]]:
        from .sweeps import pack_sweeps
        mine = frames_of(self.rank, self.world, self.num_frames)
        for group in batches(mine, self.batch):
            pts, offs = pack_sweeps([self.load(f) for f in group])
            top, front = self.encode(group, pts, offs)
            yield group, top, front


__all__ = ["init_process_group_quiet", "owner_of", "frames_of", "local_index", "shard_sizes", "batches", "interleave", "gather_frames",
           "max_over_ranks", "ShardedStream"]
