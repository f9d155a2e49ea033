"""Streaming a long run of sweeps through the device-resident encoder (BASELINE.json configs 3 and 4).

The reference encodes frame after frame in worker processes and hands each result to the next stage
(`src/data.py:486-550`, `src/utils/batch_loading.py:633-688`). Here a stream of sweeps is cut into batches; batch
k+1 is packed on the host and uploaded on its own CUDA stream while batch k is being encoded, and the results are
yielded as device tensors that stay valid until the same slot comes round again (`depth` batches later), so 4096
encoded frames of 60.8 MB never have to be resident at once.

torch is plumbing here (pinned staging, device buffers, streams); every encoding step is libmv3d_encode.so.
"""
from __future__ import annotations

import time
from typing import Callable, Iterable, Iterator, Optional, Sequence, Tuple

import numpy as np

from .encoder import ViewEncoder
from .params import FrontParams, TopParams


class _Slot:
    """One in-flight batch: its own encoder context (scratch is per context), stream, staging and outputs."""

    def __init__(self, device: int, semantics: str, strategy: str):
        import torch
        self.enc = ViewEncoder(device, semantics=semantics, strategy=strategy)
        self.stream = torch.cuda.Stream(device=device)
        self.done = torch.cuda.Event()
        self.h_pts = None   # pinned (capacity, 4) float32
        self.d_pts = None   # device twin
        self.top = None
        self.front = None
        self.frames: Sequence[int] = ()

    def stage(self, sweeps, device):
        """Pack the sweeps straight into the pinned staging buffer (one host copy); returns (total, offsets)."""
        import torch
        offs = np.zeros(len(sweeps) + 1, dtype=np.int32)
        for i, s in enumerate(sweeps):
            offs[i + 1] = offs[i] + s.shape[0]
        n = int(offs[-1])
        if self.h_pts is None or self.h_pts.shape[0] < n:
            cap = int(n * 1.25) + 1024
            self.h_pts = torch.empty((cap, 4), dtype=torch.float32, pin_memory=True)
            self.d_pts = torch.empty((cap, 4), dtype=torch.float32, device=device)
        h = self.h_pts.numpy()
        if len(sweeps) >= 4:  # parallel native copies: one Python thread packs at ~10 GB/s, a 64-sweep batch is 123 MB
            from .sweep_io import pack_sweeps_into
            pack_sweeps_into(sweeps, h)
        else:
            for i, s in enumerate(sweeps):
                h[offs[i]:offs[i + 1]] = s
        return n, offs


class SweepStream:
    """`for frames, top, front in SweepStream(...).run(frame_ids, load)`.

    * `load(frame) -> (n, 4) float32` supplies one sweep (a `.bin` reader in production);
    * `want_top=False` gives the front-view-only mode of config 3 (the bird's-eye-view gate still applies);
    * `depth` batches are in flight (three by default: one being packed on the host, one on the copy engine, one being
      encoded -- measured 24.4 k sweeps/s per GPU from host memory against 23.0 k with two, PCIe allows 28.6 k); a
      yielded `top`/`front` stays valid until the iterator has been advanced `depth - 1` more times, so consume (or copy)
      it before that.
    """

    def __init__(self, device: int, top: TopParams, front: Optional[FrontParams], batch: int = 64,
                 want_top: bool = True, depth: int = 3, semantics: str = "topfront", strategy: str = "auto"):
        import torch
        assert depth >= 2 and batch >= 1
        self.device = torch.device("cuda", device)
        self.top_p, self.front_p, self.batch, self.want_top = top, front, batch, want_top
        self.slots = [_Slot(device, semantics, strategy) for _ in range(depth)]
        # where the host thread's time went, in seconds, summed over every run() of this object: fetching the sweeps,
        # packing them into page-locked staging, queueing copy + encode, waiting for a batch to finish
        self.trace = {"load": 0.0, "pack": 0.0, "enqueue": 0.0, "wait": 0.0}

    def close(self):
        for s in self.slots:
            s.enc.close()

    @property
    def launch_count(self) -> int:
        return sum(s.enc.launch_count for s in self.slots)

    def _enqueue(self, slot: _Slot, frames: Sequence[int], load: Callable[[int], np.ndarray]):
        import torch
        t0 = time.perf_counter()
        sweeps = [load(f) for f in frames]
        t1 = time.perf_counter()
        n, offs = slot.stage(sweeps, self.device)
        t2 = time.perf_counter()
        self.trace["load"] += t1 - t0
        self.trace["pack"] += t2 - t1
        nb = len(frames)
        if self.want_top and (slot.top is None or slot.top.shape[0] < nb):
            slot.top = torch.empty((self.batch, *self.top_p.shape), dtype=torch.float32, device=self.device)
        if self.front_p is not None and (slot.front is None or slot.front.shape[0] < nb):
            slot.front = torch.empty((self.batch, *self.front_p.shape), dtype=torch.float32, device=self.device)
        # whatever the consumer queued on its (current) stream against this slot's previous result goes first
        slot.stream.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(slot.stream):
            slot.d_pts[:n].copy_(slot.h_pts[:n], non_blocking=True)
            slot.enc.encode_device(slot.d_pts[:n] if n else slot.d_pts[:0], offs, self.top_p, self.front_p,
                                   top_out=slot.top[:nb] if self.want_top else None,
                                   front_out=slot.front[:nb] if self.front_p is not None else None,
                                   want_top=self.want_top, stream=slot.stream.cuda_stream)
            slot.done.record(slot.stream)
        slot.frames = list(frames)
        self.trace["enqueue"] += time.perf_counter() - t2

    def run(self, frame_ids: Iterable[int], load: Callable[[int], np.ndarray]) -> Iterator[Tuple[Sequence[int], object, object]]:
        frame_ids = list(frame_ids)
        groups = [frame_ids[i:i + self.batch] for i in range(0, len(frame_ids), self.batch)]
        depth = len(self.slots)
        for k in range(min(depth - 1, len(groups))):  # prime the pipeline
            self._enqueue(self.slots[k % depth], groups[k], load)
        for k, group in enumerate(groups):
            nxt = k + depth - 1
            if nxt < len(groups):  # pack + upload a later batch while batch k is on the GPU
                self._enqueue(self.slots[nxt % depth], groups[nxt], load)
            slot = self.slots[k % depth]
            t0 = time.perf_counter()
            slot.done.synchronize()
            self.trace["wait"] += time.perf_counter() - t0
            nb = len(group)
            yield (slot.frames, slot.top[:nb] if self.want_top else None,
                   slot.front[:nb] if self.front_p is not None else None)


class ResidentStream:
    """BASELINE.json config 4 as SURVEY.md 8(d) specifies it: the sweeps of this rank are ALREADY in its HBM (packed
    batches), the encoded tensors go to a ring of `depth` output slots and never leave the device; the consumer sees
    each batch's slot before the ring comes round to it again. `for k, top, front in ResidentStream(...).run(batches)`
    with `batches` an iterable of `(points_dev [n, 4], offsets_host [nb + 1])`.

    Everything is enqueued on the current CUDA stream, so a consumer that enqueues its own work there (e.g.
    `ViewEncoder.frame_checksums`) is ordered before the slot is reused without any host synchronisation.
    """

    def __init__(self, device: int, top: TopParams, front: Optional[FrontParams], batch: int = 64, depth: int = 2,
                 semantics: str = "topfront", strategy: str = "auto"):
        import torch
        assert depth >= 1 and batch >= 1
        self.device = torch.device("cuda", device)
        self.top_p, self.front_p, self.batch = top, front, batch
        self.enc = ViewEncoder(device, semantics=semantics, strategy=strategy)
        self.ring = [(torch.empty((batch, *top.shape), dtype=torch.float32, device=self.device),
                      torch.empty((batch, *front.shape), dtype=torch.float32, device=self.device)
                      if front is not None else None) for _ in range(depth)]

    def close(self):
        self.enc.close()

    @property
    def launch_count(self) -> int:
        return self.enc.launch_count

    def run(self, batches) -> Iterator[Tuple[int, object, object]]:
        for k, (d_pts, offs) in enumerate(batches):
            nb = len(offs) - 1
            assert nb <= self.batch, f"batch {k} holds {nb} sweeps, the ring slots {self.batch}"
            top, front = self.ring[k % len(self.ring)]
            self.enc.encode_device(d_pts, offs, self.top_p, self.front_p, top_out=top[:nb],
                                   front_out=front[:nb] if front is not None else None)
            yield k, top[:nb], (front[:nb] if front is not None else None)


__all__ = ["SweepStream", "ResidentStream"]
