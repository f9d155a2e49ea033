"""A `KittiLoading`-shaped loader around the view encoder.

Mirrors `src/utils/batch_loading.py:550-798` for the part of a sample that comes from the LiDAR file: the same
directory layout (`<object_dir>/training/{velodyne,label_2,image_2,top_view,front_view}`), the same file listing
(sorted globs, or a split file of tags), `use_precal_view` (read `top_view/<tag>.npy` / `front_view/<tag>.npy`
written by the preprocessing run, encode when a file is missing -- the reference's try/except, `:657-666`),
`require_shuffle`, `is_testset`, `batch_size`, and `load()` returning

    testset : (tag, rgb, raw_lidar, top_view, front_view)             `:719-726`
    training: (tag, label, rgb, raw_lidar, top_view, front_view)      `:727-735`

and `None` when a test set is exhausted (`:705-706`); a training set wraps around and reshuffles (`:676-680`).

What differs: a batch of sweeps is read by native reader threads into ONE page-locked buffer
(`sweep_io.read_sweeps`) and encoded in ONE batched GPU call -- instead of one worker process per frame each
calling numpy and pycuda -- while a background thread already reads the next batch. `output="torch"` hands the
views over as CUDA tensors (nothing goes back over PCIe); `output="numpy"` returns host arrays for an unchanged TF
`feed_dict` (`src/mv3d.py:283,305`). Camera images are outside this path: pass `rgb_reader` (e.g. `cv2.imread`) to
get them, else the `rgb` slot holds `None` per frame.
"""
from __future__ import annotations

import glob
import os
import random
import threading
from typing import Callable, List, Optional, Sequence

import numpy as np

from . import sweep_io
from .params import PYTWIN_KITTI, PyTwinParams, TopParams

__all__ = ["KittiLoading", "list_frames"]


def list_frames(object_dir: str, subset: str = "training", split_file: str = ""):
    """File lists and tags exactly as `KittiLoading.__init__` builds them (`:563-584`)."""
    sub = os.path.join(object_dir, subset)
    if split_file:
        tags = [line.rstrip("\n") for line in open(split_file)]
        pick = lambda d, ext: [os.path.join(sub, d, t + ext) for t in tags]
        return {"tag": tags, "lidar": pick("velodyne", ".bin"), "label": pick("label_2", ".txt"),
                "rgb": pick("image_2", ".png"), "top": pick("top_view", ".npy"), "front": pick("front_view", ".npy")}
    pick = lambda d, ext: sorted(glob.glob(os.path.join(sub, d, "*" + ext)))
    lidar = pick("velodyne", ".bin")
    tags = [os.path.basename(p).split(".")[-2] for p in lidar]
    at = lambda d, ext: [os.path.join(sub, d, t + ext) for t in tags]
    return {"tag": tags, "lidar": lidar, "label": at("label_2", ".txt"), "rgb": at("image_2", ".png"),
            "top": at("top_view", ".npy"), "front": at("front_view", ".npy")}


class KittiLoading:
    """See the module docstring. `params`: a `PyTwinParams` (the Python encoders the reference's loader calls, default
    the `kitti` preset) or a `TopParams` (+ `front`) for the C routine's arithmetic."""

    def __init__(self, object_dir: str = ".", queue_size: int = 2, require_shuffle: bool = False,
                 is_testset: bool = True, batch_size: int = 1, use_precal_view: bool = False, split_file: str = "",
                 params=PYTWIN_KITTI, front=None, device: int = 0, output: str = "numpy",
                 rgb_reader: Optional[Callable] = None, subset: str = "training", seed: Optional[int] = None,
                 encode: Optional[Callable] = None, reader_threads: int = 0, pinned: bool = True):
        assert output in ("numpy", "torch")
        self.object_dir, self.is_testset, self.batch_size = object_dir, is_testset, int(batch_size)
        self.require_shuffle = require_shuffle if not is_testset else False   # `:557`
        self.use_precal_view, self.output = use_precal_view, output
        self.params, self.front, self.device = params, front, device
        self.rgb_reader, self.reader_threads, self.pinned = rgb_reader, reader_threads, pinned
        self.files = list_frames(object_dir, subset, split_file)
        self.dataset_size = len(self.files["lidar"])
        self.alreay_extract_data = 0       # (sic) the reference's counter name
        self.cur_frame_info = ""
        self._rng = random.Random(seed)
        self._order = list(range(self.dataset_size))
        if self.require_shuffle:
            self._rng.shuffle(self._order)
        self._cursor = 0
        self._encode = encode              # injected in the CPU tests; None = the GPU encoder, created on first use
        self._encoder = None
        self._depth = max(1, int(queue_size))
        self._pending: List = []           # [(thread, slot)] batches being read ahead
        if isinstance(params, PyTwinParams):
            self.top_shape, self.front_shape = params.top_shape, params.front_shape
        else:
            self.top_shape, self.front_shape = params.shape, (front.shape if front is not None else None)

    # -- reference surface ----------------------------------------------------------------------------------
    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __len__(self):
        return self.dataset_size

    def get_shape(self):
        return self.top_shape, self.front_shape, None

    def get_frame_info(self):
        return self.cur_frame_info

    def close(self):
        for th, _ in self._pending:
            th.join()
        self._pending = []
        if self._encoder is not None:
            self._encoder.close()
            self._encoder = None

    # -- batches --------------------------------------------------------------------------------------------
    def _next_indices(self) -> Optional[List[int]]:
        if self.dataset_size == 0:
            return None
        if self.is_testset:
            if self._cursor >= self.dataset_size:
                return None
            idx = self._order[self._cursor:self._cursor + self.batch_size]
            self._cursor += len(idx)
            return idx
        idx = []
        while len(idx) < self.batch_size:                      # wrap around, reshuffling (`:676-680`)
            if self._cursor >= self.dataset_size:
                self._cursor = 0
                if self.require_shuffle:
                    self._rng.shuffle(self._order)
            idx.append(self._order[self._cursor])
            self._cursor += 1
        return idx

    def _read(self, slot: dict) -> None:
        try:
            idx = slot["idx"]
            slot["points"], slot["offsets"] = sweep_io.read_sweeps([self.files["lidar"][i] for i in idx],
                                                                   threads=self.reader_threads, pinned=self.pinned)
            if not self.is_testset:
                slot["label"] = [open(self.files["label"][i]).readlines() for i in idx]   # `:670`
            slot["rgb"] = [self.rgb_reader(self.files["rgb"][i]) if self.rgb_reader else None for i in idx]
            if self.use_precal_view:
                slot["top_files"] = [self.files["top"][i] if os.path.exists(self.files["top"][i]) else None for i in idx]
                slot["front_files"] = [self.files["front"][i] if os.path.exists(self.files["front"][i]) else None
                                       for i in idx]
        except Exception as e:  # surfaced by load() on the consumer's thread
            slot["error"] = e

    def _schedule(self) -> None:
        while len(self._pending) < self._depth:
            idx = self._next_indices()
            if idx is None:
                return
            slot = {"idx": idx}
            th = threading.Thread(target=self._read, args=(slot,), daemon=True)
            th.start()
            self._pending.append((th, slot))

    def _encode_batch(self, points, offsets, want_top=True, want_front=True):
        if self._encode is not None:
            return self._encode(points, offsets, want_top, want_front)
        pytwin = isinstance(self.params, PyTwinParams)
        if self._encoder is None:
            if pytwin:
                from .pytwin import PyTwinEncoder
                self._encoder = PyTwinEncoder(self.device)
            else:
                from .encoder import ViewEncoder
                self._encoder = ViewEncoder(self.device, semantics="topfront" if self.front is not None else "top")
        want_front = want_front and (pytwin or self.front is not None)
        if self.output == "torch":
            import torch
            d = torch.from_numpy(points).to(f"cuda:{self.device}", non_blocking=True)
            if pytwin:
                return self._encoder.encode_device(d, offsets, self.params, want_top=want_top, want_front=want_front)
            return self._encoder.encode_device(d, offsets, self.params, self.front if want_front else None,
                                               want_top=want_top)
        if pytwin:
            return self._encoder.encode_host(points, offsets, self.params, want_top=want_top, want_front=want_front)
        return self._encoder.encode_host(points, offsets, self.params, self.front if want_front else None,
                                         want_top=want_top)

    def _views(self, slot):
        points, offsets = slot["points"], slot["offsets"]
        n = len(slot["idx"])
        tops, fronts = slot.get("top_files"), slot.get("front_files")
        if not self.use_precal_view or (all(t is None for t in tops) and all(f is None for f in fronts)):
            return self._encode_batch(points, offsets)
        # precomputed views where they exist, the encoder for the frames that miss one (`:657-666`)
        miss_top = [k for k in range(n) if tops[k] is None]
        miss_front = [k for k in range(n) if fronts[k] is None] if self.front_shape is not None else []
        todo = sorted(set(miss_top) | set(miss_front))
        top = np.empty((n, *self.top_shape), np.float32)
        front = np.empty((n, *self.front_shape), np.float32) if self.front_shape is not None else None
        for k in range(n):
            if tops[k] is not None:
                sweep_io.load_view(tops[k], out=top[k])
            if front is not None and fronts[k] is not None:
                sweep_io.load_view(fronts[k], out=front[k])
        if todo:
            sub_off = np.zeros(len(todo) + 1, np.int32)
            parts = []
            for j, k in enumerate(todo):
                parts.append(points[offsets[k]:offsets[k + 1]])
                sub_off[j + 1] = sub_off[j] + parts[-1].shape[0]
            et, ef = self._encode_batch(np.ascontiguousarray(np.concatenate(parts)), sub_off)
            to_np = lambda t: t if isinstance(t, np.ndarray) else t.cpu().numpy()
            for j, k in enumerate(todo):
                if k in miss_top:
                    top[k] = to_np(et[j])
                if front is not None and k in miss_front:
                    front[k] = to_np(ef[j])
        if self.output == "torch":
            import torch
            dev = f"cuda:{self.device}"
            return torch.from_numpy(top).to(dev), (torch.from_numpy(front).to(dev) if front is not None else None)
        return top, front

    def load(self):
        """One batch, or None when a test set is exhausted."""
        self._schedule()
        if not self._pending:
            return None
        th, slot = self._pending.pop(0)
        th.join()
        self._schedule()                      # the next batch is read while this one is encoded
        if "error" in slot:
            raise slot["error"]
        top, front = self._views(slot)
        idx, offsets, points = slot["idx"], slot["offsets"], slot["points"]
        tag = np.array([self.files["tag"][i] for i in idx])
        raw_lidar = [points[offsets[k]:offsets[k + 1]] for k in range(len(idx))]
        self.alreay_extract_data += len(idx)
        self.cur_frame_info = str(tag[-1])
        if self.is_testset:
            return tag, slot["rgb"], raw_lidar, top, front
        return tag, slot["label"], slot["rgb"], raw_lidar, top, front

    def __iter__(self):
        while True:
            batch = self.load()
            if batch is None:
                return
            yield batch
