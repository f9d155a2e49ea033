"""Host-side mirror of the reference's empty-anchor filter, bound to libmv3d_encode.so (include/mv3d_anchors.h).

`remove_empty_anchor(view, anchors, limit)` keeps the name, arguments and return value of
`src/net/utility/remove_empty_box.py:25-47`: a `(W, H, C)` float32 view and `(N, 4)` int32 anchors
`(y1, x1, y2, x2)` in, the ascending indices of the anchors whose box sum exceeds `limit` out (what
`np.where(...)[0]` returns), so `src/mv3d.py:280,343,1139,1460` can call it unchanged.

`AnchorFilter` is the form the path is built for: the anchor set is uploaded once, and the views are the tensors the
encoder just wrote in device memory (no upload of 52 MB per frame, a whole batch of frames per call). `make_anchors`
rebuilds the reference's anchor set (`src/net/rpn_target_op.py:86-116`, bases `src/mv3d.py:186-191`) for benches and
tests. All filtering arithmetic is in the library.
"""
from __future__ import annotations

import ctypes
import math

import numpy as np

from . import _lib

__all__ = ["remove_empty_anchor", "AnchorFilter", "make_anchors", "MV3D_BASES"]

# The four top-view anchor shapes of src/mv3d.py:186-191 (x1, y1, x2, y2 around a feature-map position).
MV3D_BASES = np.array([[4.5, 2.5, 10.5, 12.5], [2.5, 4.5, 12.5, 10.5], [-0.5, -12, 15.5, 27], [-12, -0.5, 27, 15.5]])


def make_anchors(top_shape, stride: int = 8, bases=MV3D_BASES) -> np.ndarray:
    """int32 `(H*W*len(bases), 4)`: every base shifted to every feature-map position (`make_anchors`,
    src/net/rpn_target_op.py:86-103; feature shape = ceil(top_shape / stride), src/mv3d.py:68-69)."""
    fh, fw = math.ceil(top_shape[0] / stride), math.ceil(top_shape[1] / stride)
    sx, sy = np.meshgrid(np.arange(fw) * stride, np.arange(fh) * stride)
    shifts = np.stack([sx.ravel(), sy.ravel(), sx.ravel(), sy.ravel()], axis=1)
    anchors = np.asarray(bases, dtype=np.float64)[None, :, :] + shifts[:, None, :]
    return anchors.reshape(-1, 4).astype(np.int32)


class AnchorFilter:
    """Filters one resident anchor set against device-resident views. `owner`: a `ViewEncoder` / `PyTwinEncoder`
    whose context (and GPU) to share, or a device index for a context of its own."""

    def __init__(self, owner=0, anchors=None, mode: str = "auto"):
        self._lib = _lib.load_library()
        self._own = None
        if isinstance(owner, int):
            from .encoder import ViewEncoder
            self._own = owner = ViewEncoder(owner)
        self._owner = owner              # keeps the context alive
        self.mode = {"auto": _lib.ANCHORS_AUTO, "exact": _lib.ANCHORS_EXACT}[mode]
        self.n = 0
        if anchors is not None:
            self.set_anchors(anchors)

    @property
    def _ctx(self):
        return self._owner._ctx

    def close(self):
        if self._own is not None:
            self._own.close()
            self._own = None

    def set_anchors(self, anchors) -> None:
        anchors = np.ascontiguousarray(anchors, dtype=np.int32).reshape(-1, 4)
        if self._lib.mv3d_ctx_set_anchors(self._ctx, anchors.ctypes.data, anchors.shape[0]) != 0:
            raise RuntimeError("mv3d_ctx_set_anchors: " + _lib.last_error())
        self.n = anchors.shape[0]

    def filter_device(self, views, limit: float = 0.0, want_totals: bool = False, stream=None):
        """`views`: CUDA float32 `[frames, W, H, C]` (or `[W, H, C]`). Returns `(keep, count[, totals])`: int32 CUDA
        tensors `[frames, n]` / `[frames]`; row f of `keep` starts with the `count[f]` kept anchor indices in
        ascending order. Only enqueues work on `stream` (default: torch's current stream)."""
        import torch
        assert views.is_cuda and views.dtype == torch.float32 and views.is_contiguous()
        if views.dim() == 3:
            views = views.unsqueeze(0)
        frames, xn, yn, planes = views.shape
        keep = torch.empty((frames, self.n), dtype=torch.int32, device=views.device)
        count = torch.empty((frames,), dtype=torch.int32, device=views.device)
        totals = torch.empty((frames, self.n), dtype=torch.float32, device=views.device) if want_totals else None
        if stream is None:
            stream = torch.cuda.current_stream(views.device).cuda_stream
        rc = self._lib.mv3d_remove_empty_anchor_dev(
            self._ctx, views.data_ptr(), frames, xn, yn, planes, float(limit), self.mode, keep.data_ptr(),
            count.data_ptr(), totals.data_ptr() if want_totals else None, ctypes.c_void_p(stream))
        if rc != 0:
            raise RuntimeError("mv3d_remove_empty_anchor_dev: " + _lib.last_error())
        return (keep, count, totals) if want_totals else (keep, count)

    def encode_and_filter(self, points, offsets, params, front=None, limit: float = 0.0, top_out=None, front_out=None,
                          want_front: bool = True, stream=None):
        """Encode `len(offsets)-1` device-resident sweeps with the owning encoder and filter the views it writes, in
        one library call: the kernel that builds a view also reports which cells it touched, so the filter does not
        read the view again. `params`: `TopParams` (+ optional `front`: `FrontParams`) for a `ViewEncoder`, a
        `PyTwinParams` for a `PyTwinEncoder`. Returns `(top, front, keep, count)` -- the same tensors `encode_device`
        followed by `filter_device` would give."""
        import torch
        from .params import PyTwinParams
        assert points.is_cuda and points.dtype == torch.float32 and points.is_contiguous()
        offs = np.ascontiguousarray(offsets, dtype=np.int32)
        batch = len(offs) - 1
        offs_p = offs.ctypes.data_as(ctypes.POINTER(ctypes.c_int))
        dev = points.device
        keep = torch.empty((batch, self.n), dtype=torch.int32, device=dev)
        count = torch.empty((batch,), dtype=torch.int32, device=dev)
        if stream is None:
            stream = torch.cuda.current_stream(dev).cuda_stream
        if isinstance(params, PyTwinParams):
            from .pytwin import _params_c
            if top_out is None:
                top_out = torch.empty((batch, *params.top_shape), dtype=torch.float32, device=dev)
            if want_front and front_out is None:
                front_out = torch.empty((batch, *params.front_shape), dtype=torch.float32, device=dev)
            pc = _params_c(params)
            rc = self._lib.mv3d_pytwin_encode_filter_batch_dev(
                self._ctx, points.data_ptr(), offs_p, batch, top_out.data_ptr(),
                front_out.data_ptr() if want_front else None, ctypes.byref(pc), float(limit), self.mode,
                keep.data_ptr(), count.data_ptr(), ctypes.c_void_p(stream))
            name = "mv3d_pytwin_encode_filter_batch_dev"
        else:
            from .encoder import _front_c, _top_c
            if top_out is None:
                top_out = torch.empty((batch, *params.shape), dtype=torch.float32, device=dev)
            if front is not None and front_out is None:
                front_out = torch.empty((batch, *front.shape), dtype=torch.float32, device=dev)
            rc = self._lib.mv3d_encode_filter_batch_dev(
                self._ctx, points.data_ptr(), offs_p, batch, top_out.data_ptr(),
                front_out.data_ptr() if front is not None else None, ctypes.byref(_top_c(params)),
                ctypes.byref(_front_c(front)) if front is not None else None, self._owner.semantics,
                self._owner.strategy, float(limit), self.mode, keep.data_ptr(), count.data_ptr(),
                ctypes.c_void_p(stream))
            name = "mv3d_encode_filter_batch_dev"
        if rc != 0:
            raise RuntimeError(name + ": " + _lib.last_error())
        return top_out, front_out, keep, count

    def filter_host(self, view: np.ndarray, anchors, limit: float = 0.0) -> np.ndarray:
        """The reference's call: numpy in, kept indices (int64, ascending) out."""
        view = np.ascontiguousarray(view, dtype=np.float32)
        anchors = np.ascontiguousarray(anchors, dtype=np.int32).reshape(-1, 4)
        assert view.ndim == 3, "view must be (W, H, C)"
        keep = np.empty(anchors.shape[0], dtype=np.int32)
        count = ctypes.c_int(0)
        rc = self._lib.mv3d_remove_empty_anchor_host(
            self._ctx, view.ctypes.data, view.shape[0], view.shape[1], view.shape[2], anchors.ctypes.data,
            anchors.shape[0], float(limit), self.mode, keep.ctypes.data, ctypes.byref(count))
        if rc != 0:
            raise RuntimeError("mv3d_remove_empty_anchor_host: " + _lib.last_error())
        self.n = anchors.shape[0]
        return keep[: count.value].astype(np.int64)


_default = None


def remove_empty_anchor(view, anchors, limit) -> np.ndarray:
    """`src/net/utility/remove_empty_box.py:25` on the GPU."""
    global _default
    if _default is None:
        import os
        _default = AnchorFilter(int(os.environ.get("MV3D_DEVICE", "0")))
    return _default.filter_host(view, anchors, limit)
