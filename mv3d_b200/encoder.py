"""Device-resident, batched view encoder: thin Python over the additive C ABI (include/mv3d_encode.h).

PyTorch is plumbing here (device memory, streams, torch.distributed); every computation happens in
libmv3d_encode.so. Tensors are handed over by `data_ptr()`.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _lib
from .params import FrontParams, TopParams


def _top_c(top: TopParams) -> _lib.TopParamsC:
    return _lib.TopParamsC(*top.floats, top.xn, top.yn, top.zn, top.flags)


def _front_c(front: FrontParams) -> _lib.FrontParamsC:
    return _lib.FrontParamsC(front.phi_min, front.phi_max, front.theta_min, front.theta_max,
                             front.phi_div, front.theta_div, float(front.cn), float(front.rn), float(front.fn))


class ViewEncoder:
    """One per GPU. `semantics`: "top" = createTopMaps arithmetic (truncating cell index),
    "topfront" = createTopAndFrontMaps arithmetic (floor). `strategy`: "auto" | "tile" | "atomic"."""

    def __init__(self, device: int = 0, semantics: str = "topfront", strategy: str = "auto"):
        self._lib = _lib.load_library()
        self._ctx = ctypes.c_void_p()
        if self._lib.mv3d_ctx_create(int(device), ctypes.byref(self._ctx)) != 0:
            raise RuntimeError("mv3d_ctx_create: " + _lib.last_error())
        self.device = int(device)
        self.semantics = {"top": _lib.SEM_TOP, "topfront": _lib.SEM_TOPFRONT}[semantics]
        self.strategy = _lib.STRATEGIES[strategy]

    def close(self):
        if getattr(self, "_ctx", None):
            self._lib.mv3d_ctx_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launch_count(self) -> int:
        return int(self._lib.mv3d_ctx_launch_count(self._ctx))

    def set_profiling(self, enable: bool) -> None:
        """Record CUDA events around every kernel (read back with `kernel_times`)."""
        self._lib.mv3d_ctx_set_profiling(self._ctx, int(bool(enable)))

    def kernel_times(self, reset: bool = False) -> dict:
        """{kernel name: (accumulated device ms, launches)}; waits for the recorded events."""
        n = self._lib.mv3d_kernel_count()
        ms = (ctypes.c_double * n)()
        cnt = (ctypes.c_longlong * n)()
        if self._lib.mv3d_ctx_kernel_times(self._ctx, ms, cnt, int(reset)) != 0:
            raise RuntimeError("mv3d_ctx_kernel_times: " + _lib.last_error())
        return {self._lib.mv3d_kernel_name(k).decode(): (float(ms[k]), int(cnt[k])) for k in range(n)}

    def transfer_bytes(self, reset: bool = False):
        """(host->device, device->host) bytes moved by `encode_host` since the last reset."""
        h2d, d2h = ctypes.c_ulonglong(), ctypes.c_ulonglong()
        if self._lib.mv3d_ctx_transfer_bytes(self._ctx, ctypes.byref(h2d), ctypes.byref(d2h), int(reset)) != 0:
            raise RuntimeError("mv3d_ctx_transfer_bytes: " + _lib.last_error())
        return int(h2d.value), int(d2h.value)

    def resolved_strategy(self, top: TopParams) -> str:
        code = self._lib.mv3d_resolve_strategy(self._ctx, ctypes.byref(_top_c(top)), self.strategy)
        return {v: k for k, v in _lib.STRATEGIES.items()}[code]

    # -- device-resident path ---------------------------------------------------------------------
    def encode_device(self, points, offsets, top: TopParams, front: FrontParams | None = None,
                      top_out=None, front_out=None, want_top: bool = True, stream=None):
        """Encode `batch = len(offsets)-1` sweeps held in one CUDA tensor `points` (total, 4) float32.

        Returns `(top_out, front_out)`: float32 CUDA tensors `[batch, Xn, Yn, Zn+2]` / `[batch, R, C, F]`
        (None for a view that was not requested). Only enqueues work on `stream` (default: torch's current).
        """
        import torch
        assert points.is_cuda and points.dtype == torch.float32 and points.is_contiguous()
        offs = np.ascontiguousarray(offsets, dtype=np.int32)
        batch = len(offs) - 1
        _check_offsets(offs, points.numel() // 4)
        if want_top and top_out is None:
            top_out = torch.empty((batch, *top.shape), dtype=torch.float32, device=points.device)
        if front is not None and front_out is None:
            front_out = torch.empty((batch, *front.shape), dtype=torch.float32, device=points.device)
        for name, out, shape in (("top_out", top_out if want_top else None, top.shape),
                                 ("front_out", front_out, front.shape if front is not None else None)):
            if out is not None:  # a wrong buffer is an out-of-bounds device write, not an exception: refuse it here
                if not (out.is_cuda and out.device == points.device and out.dtype == torch.float32 and out.is_contiguous()):
                    raise ValueError(f"{name} must be a contiguous float32 CUDA tensor on {points.device}")
                if tuple(out.shape) != (batch, *shape):
                    raise ValueError(f"{name} has shape {tuple(out.shape)}, expected {(batch, *shape)}")
        if stream is None:
            stream = torch.cuda.current_stream(points.device).cuda_stream
        rc = self._lib.mv3d_encode_batch_dev(
            self._ctx, points.data_ptr(), offs.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), batch,
            top_out.data_ptr() if (want_top and top_out is not None) else None,
            front_out.data_ptr() if front_out is not None else None,
            ctypes.byref(_top_c(top)), ctypes.byref(_front_c(front)) if front is not None else None,
            self.semantics, self.strategy, ctypes.c_void_p(stream))
        if rc != 0:
            raise RuntimeError("mv3d_encode_batch_dev: " + _lib.last_error())
        return (top_out if want_top else None), front_out

    def frame_checksums(self, views, out=None, stream=None):
        """Sum of the 32-bit words of every frame of a CUDA float32 tensor `[frames, ...]`, modulo 2^64, as an int64
        CUDA tensor (the bit pattern of the unsigned sum): the on-device consumer of a streamed run."""
        import torch
        assert views.is_cuda and views.dtype == torch.float32 and views.is_contiguous()
        frames = views.shape[0]
        if out is None:
            out = torch.empty(frames, dtype=torch.int64, device=views.device)
        if stream is None:
            stream = torch.cuda.current_stream(views.device).cuda_stream
        per = views.numel() // frames if frames else 0
        rc = self._lib.mv3d_frame_checksums_dev(self._ctx, views.data_ptr(), frames, per, out.data_ptr(),
                                                ctypes.c_void_p(stream))
        if rc != 0:
            raise RuntimeError("mv3d_frame_checksums_dev: " + _lib.last_error())
        return out

    def debug_point_keys(self, points, top: TopParams, front: FrontParams | None = None):
        """Test hook: per-point (cell, front pixel, front pixel by the reference's own expressions -- no fast path)
        as int32 CUDA tensors."""
        import torch
        n = points.shape[0]
        cell = torch.empty(n, dtype=torch.int32, device=points.device)
        pix = torch.empty(n, dtype=torch.int32, device=points.device) if front is not None else None
        pix_exact = torch.empty(n, dtype=torch.int32, device=points.device) if front is not None else None
        rc = self._lib.mv3d_debug_point_keys(
            self._ctx, points.data_ptr(), n, ctypes.byref(_top_c(top)),
            ctypes.byref(_front_c(front)) if front is not None else None, self.semantics, cell.data_ptr(),
            pix.data_ptr() if pix is not None else None, pix_exact.data_ptr() if pix_exact is not None else None,
            ctypes.c_void_p(torch.cuda.current_stream(points.device).cuda_stream))
        if rc != 0:
            raise RuntimeError("mv3d_debug_point_keys: " + _lib.last_error())
        return cell, pix, pix_exact

    # -- host buffers (H2D + encode + D2H, pipelined inside the library) -----------------------------
    def encode_host(self, points: np.ndarray, offsets, top: TopParams, front: FrontParams | None = None,
                    top_out: np.ndarray | None = None, front_out: np.ndarray | None = None, want_top: bool = True):
        """Same, for numpy buffers. Pass arrays from `pinned_empty` to get direct DMA."""
        offs = np.ascontiguousarray(offsets, dtype=np.int32)
        batch = len(offs) - 1
        assert points.dtype == np.float32 and points.flags["C_CONTIGUOUS"]
        _check_offsets(offs, points.size // 4)
        if want_top and top_out is None:
            top_out = np.empty((batch, *top.shape), dtype=np.float32)
        if front is not None and front_out is None:
            front_out = np.empty((batch, *front.shape), dtype=np.float32)
        for name, out, shape in (("top_out", top_out if want_top else None, top.shape),
                                 ("front_out", front_out, front.shape if front is not None else None)):
            if out is not None:
                if not (isinstance(out, np.ndarray) and out.dtype == np.float32 and out.flags["C_CONTIGUOUS"] and out.flags["WRITEABLE"]):
                    raise ValueError(f"{name} must be a writable C-contiguous float32 numpy array")
                if out.shape != (batch, *shape):
                    raise ValueError(f"{name} has shape {out.shape}, expected {(batch, *shape)}")
        rc = self._lib.mv3d_encode_batch_host(
            self._ctx, points.ctypes.data, offs.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), batch,
            top_out.ctypes.data if (want_top and top_out is not None) else None,
            front_out.ctypes.data if front_out is not None else None,
            ctypes.byref(_top_c(top)), ctypes.byref(_front_c(front)) if front is not None else None,
            self.semantics, self.strategy)
        if rc != 0:
            raise RuntimeError("mv3d_encode_batch_host: " + _lib.last_error())
        return (top_out if want_top else None), front_out


def _check_offsets(offs: np.ndarray, n_points: int) -> None:
    """The library trusts `offsets` against the point buffer; a mismatch would be an out-of-bounds read."""
    if offs.ndim != 1 or offs.size < 1:
        raise ValueError("offsets must be a 1-d array of batch + 1 entries")
    # order / sign errors are the library's to report (check_offsets in mv3d_abi.cu); what it cannot see is the buffer
    if int(offs.max()) > n_points:
        raise ValueError(f"offsets reach point {int(offs.max())}, the point buffer holds {n_points}")


def pinned_empty(shape, dtype=np.float32) -> np.ndarray:
    """A numpy array on page-locked host memory (mv3d_host_alloc), so copies to/from the GPU are direct DMA.
    The allocation is released when the last array viewing it is garbage-collected."""
    import weakref
    dtype = np.dtype(dtype)
    count = int(np.prod(shape, dtype=np.int64))
    nbytes = max(count * dtype.itemsize, 1)
    lib = _lib.load_library()
    ptr = lib.mv3d_host_alloc(nbytes)
    if not ptr:
        raise MemoryError("mv3d_host_alloc: " + _lib.last_error())
    buf = (ctypes.c_char * nbytes).from_address(ptr)
    weakref.finalize(buf, lib.mv3d_host_free, ptr)
    return np.frombuffer(buf, dtype=dtype, count=count).reshape(shape)
