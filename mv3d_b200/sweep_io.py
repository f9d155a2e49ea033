"""On-disk formats either side of the view encoder, bound to libmv3d_encode.so (include/mv3d_io.h).

What the reference does with numpy one file at a time -- `np.fromfile(f, np.float32).reshape((-1, 4))`
(`src/utils/batch_loading.py:656`, `src/raw_data.py:137-139`), `np.save(path, view)`
(`src/preprocess_for_mv3d.py:36-37,56-57`), `np.savez_compressed(path, top_view=top)` (`src/data.py:521`) -- done
for a whole batch by native reader / writer threads, straight into and out of the page-locked buffers the encoder
uploads from and downloads to. File contents are the reference's formats: `.npy` files are byte-identical to
numpy's, `.npz` archives load with `np.load(path)['top_view']`.
"""
from __future__ import annotations

import ctypes
import os
from typing import Sequence

import numpy as np

from . import _lib
from .encoder import pinned_empty

__all__ = ["read_sweep", "read_sweeps", "pack_sweeps_into", "save", "save_batch", "load_view", "savez_compressed",
           "savez_compressed_batch", "npy_path", "npz_path"]


def _paths(paths: Sequence[str]):
    enc = [os.fsencode(p) for p in paths]
    return (ctypes.c_char_p * len(enc))(*enc), enc


def _shape(shape):
    return (ctypes.c_longlong * max(len(shape), 1))(*shape), len(shape)


def _f32(a: np.ndarray) -> np.ndarray:
    if a.dtype != np.float32 or not a.flags["C_CONTIGUOUS"]:
        raise TypeError("views are written from C-contiguous float32 arrays")
    return a


def npy_path(path: str) -> str:
    """np.save appends '.npy' unless the name already ends with it."""
    return path if path.endswith(".npy") else path + ".npy"


def npz_path(path: str) -> str:
    """np.savez_compressed appends '.npz' unless the name already ends with it (src/data.py:518-521 therefore writes
    '00013.npy.npz', which src/utils/batch_loading.py:51 reads back)."""
    return path if path.endswith(".npz") else path + ".npz"


def sweep_sizes(paths: Sequence[str]) -> np.ndarray:
    """Points per `.bin` file; raises where numpy's reshape((-1, 4)) would."""
    lib = _lib.load_library()
    arr, _keep = _paths(paths)
    out = np.zeros(len(paths), dtype=np.int64)
    if lib.mv3d_io_stat_sweeps(arr, len(paths), out.ctypes.data) != 0:
        raise ValueError(_lib.last_error())
    return out


def read_sweeps(paths: Sequence[str], threads: int = 0, pinned: bool = True, staging: "SweepStaging | None" = None):
    """Read a batch of KITTI `.bin` sweeps into ONE packed `(total, 4)` float32 array (page-locked by default) and
    return `(points, offsets)` -- what `pack_sweeps` makes of a list of arrays, and what `ViewEncoder.encode_host` /
    a single `cudaMemcpyAsync` want. A caller that reads batch after batch passes a `SweepStaging`: the page-locked
    buffer is then reused (allocating one per batch costs more than reading the files), and `points` is a view of it that
    stays valid until the same staging object is used again."""
    lib = _lib.load_library()
    sizes = sweep_sizes(paths)
    offsets = np.zeros(len(paths) + 1, dtype=np.int64)
    np.cumsum(sizes, out=offsets[1:])
    total = int(offsets[-1])
    if staging is not None:
        points = staging.take(total)
    else:
        points = pinned_empty((total, 4), np.float32) if pinned else np.empty((total, 4), np.float32)
    arr, _keep = _paths(paths)
    if lib.mv3d_io_read_sweeps(arr, len(paths), points.ctypes.data, offsets.ctypes.data, int(threads)) != 0:
        raise OSError(_lib.last_error())
    return points, offsets.astype(np.int32)


class SweepStaging:
    """A grow-only page-locked staging buffer for `read_sweeps`: `take(n)` returns a `(n, 4)` float32 view of it."""

    def __init__(self):
        self._buf = None  # allocated on first use: a run that finds nothing to read needs no GPU driver

    def take(self, points: int) -> np.ndarray:
        if self._buf is None or points > self._buf.shape[0]:
            self._buf = pinned_empty((points + points // 4 + 1024, 4), np.float32)
        return self._buf[:points]


def pack_sweeps_into(sweeps: Sequence[np.ndarray], dst: np.ndarray, threads: int = 0) -> np.ndarray:
    """Copy in-memory sweeps (`(n_i, 4)` float32 arrays) back to back into `dst` (typically a page-locked staging
    buffer) with parallel native copies; returns the int32 offsets (`len(sweeps) + 1`)."""
    import ctypes
    lib = _lib.load_library()
    sweeps = [np.ascontiguousarray(s, dtype=np.float32) for s in sweeps]
    offsets = np.zeros(len(sweeps) + 1, dtype=np.int64)
    for i, s in enumerate(sweeps):
        assert s.ndim == 2 and s.shape[1] == 4, "a sweep is (n, 4) float32"
        offsets[i + 1] = offsets[i] + s.shape[0]
    assert dst.dtype == np.float32 and dst.flags["C_CONTIGUOUS"] and dst.size >= 4 * int(offsets[-1])
    ptrs = (ctypes.c_void_p * max(len(sweeps), 1))(*[s.ctypes.data for s in sweeps])
    if lib.mv3d_io_pack_sweeps(ptrs, len(sweeps), dst.ctypes.data, offsets.ctypes.data, int(threads)) != 0:
        raise ValueError(_lib.last_error())
    return offsets.astype(np.int32)


def read_sweep(path: str) -> np.ndarray:
    """`np.fromfile(path, dtype=np.float32).reshape((-1, 4))`."""
    points, _ = read_sweeps([path], threads=1, pinned=False)
    return points


def save(path: str, view: np.ndarray) -> None:
    """`np.save(path, view)` for a float32 view."""
    lib = _lib.load_library()
    view = _f32(np.asarray(view))
    shape, ndim = _shape(view.shape)
    if lib.mv3d_io_write_npy(os.fsencode(npy_path(path)), view.ctypes.data, shape, ndim) != 0:
        raise OSError(_lib.last_error())


def save_batch(paths: Sequence[str], views: np.ndarray, threads: int = 0) -> None:
    """`for p, v in zip(paths, views): np.save(p, v)` with parallel writers; `views` is one `[n, ...]` array."""
    lib = _lib.load_library()
    views = _f32(np.asarray(views))
    if views.shape[0] != len(paths):
        raise ValueError(f"{len(paths)} paths for {views.shape[0]} views")
    arr, _keep = _paths([npy_path(p) for p in paths])
    shape, ndim = _shape(views.shape[1:])
    stride = int(np.prod(views.shape[1:], dtype=np.int64))
    if lib.mv3d_io_write_npy_batch(arr, len(paths), views.ctypes.data, stride, shape, ndim, int(threads)) != 0:
        raise OSError(_lib.last_error())


def load_view(path: str, out: np.ndarray | None = None, pinned: bool = False) -> np.ndarray:
    """`np.load(path)` of a float32 `.npy` view (src/utils/batch_loading.py:659-667), optionally into `out` or into
    page-locked memory."""
    lib = _lib.load_library()
    shape = (ctypes.c_longlong * 8)()
    ndim, off = ctypes.c_int(0), ctypes.c_longlong(0)
    if lib.mv3d_io_npy_info(os.fsencode(path), shape, ctypes.byref(ndim), ctypes.byref(off)) != 0:
        raise OSError(_lib.last_error())
    shp = tuple(int(shape[i]) for i in range(ndim.value))
    if out is None:
        out = pinned_empty(shp, np.float32) if pinned else np.empty(shp, np.float32)
    elif out.dtype != np.float32 or not out.flags["C_CONTIGUOUS"] or out.size < int(np.prod(shp, dtype=np.int64)):
        raise ValueError("out must be a C-contiguous float32 array with room for the view")
    if lib.mv3d_io_read_npy(os.fsencode(path), out.ctypes.data, out.size) != 0:
        raise OSError(_lib.last_error())
    return out.reshape(shp) if out.shape != shp else out


def savez_compressed(path: str, level: int = 6, **arrays) -> None:
    """`np.savez_compressed(path, top_view=top)` for exactly one float32 array (the reference's use, src/data.py:521)."""
    if len(arrays) != 1:
        raise ValueError("one array per archive, e.g. savez_compressed(path, top_view=top)")
    (member, view), = arrays.items()
    lib = _lib.load_library()
    view = _f32(np.asarray(view))
    shape, ndim = _shape(view.shape)
    if lib.mv3d_io_write_npz_compressed(os.fsencode(npz_path(path)), member.encode(), view.ctypes.data, shape, ndim,
                                        int(level)) != 0:
        raise OSError(_lib.last_error())


def savez_compressed_batch(paths: Sequence[str], views: np.ndarray, member: str = "top_view", level: int = 6,
                           threads: int = 0) -> None:
    """One archive per view of `views[n, ...]`, deflated by parallel workers (the zlib pass is the expensive step of
    the reference's `generate_top_view`, src/data.py:487-524)."""
    lib = _lib.load_library()
    views = _f32(np.asarray(views))
    if views.shape[0] != len(paths):
        raise ValueError(f"{len(paths)} paths for {views.shape[0]} views")
    arr, _keep = _paths([npz_path(p) for p in paths])
    shape, ndim = _shape(views.shape[1:])
    stride = int(np.prod(views.shape[1:], dtype=np.int64))
    if lib.mv3d_io_write_npz_compressed_batch(arr, len(paths), member.encode(), views.ctypes.data, stride, shape, ndim,
                                              int(level), int(threads)) != 0:
        raise OSError(_lib.last_error())
