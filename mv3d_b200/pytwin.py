"""Host-side mirror of the reference's PYTHON view encoders, bound to libmv3d_encode.so (Python-twin mode).

`lidar_to_top(lidar)` and `lidar_to_front_fast(lidar)` keep the names, argument and return conventions of
`src/data.py:296-367` and `src/data.py:113-180` (`Preprocess.lidar_to_front_fast`): an `(N, 4)` float32 numpy array
in, a float32 numpy array out -- `[Xn][Yn][Zn+2]`, already flipped as the reference writes it, resp.
`[FRONT_WIDTH][FRONT_HEIGHT][3]` -- so `src/preprocess_for_mv3d.py:30-32` and the loaders
(`src/utils/batch_loading.py:659-672`) can call them unchanged. The reference reads its parameters from module-level
constants (`src/config.py`); here they come from a `PyTwinParams` (default: the reference's `kitti` preset).

Deviation, on purpose: the reference's `lidar_to_front_fast` raises ValueError when no point survives the crop and its
`lidar_to_top` breaks on exactly one surviving point (numpy `squeeze`); these functions return the all-zero / the
correct tensor instead.

`PyTwinEncoder` is the device-resident, batched form (torch tensors in and out). All arithmetic is in the library.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _lib
from .params import PYTWIN_KITTI, PyTwinParams

__all__ = ["lidar_to_top", "lidar_to_front_fast", "lidar_to_top_cuda", "lidar_to_front_cuda", "Preprocess", "PyTwinEncoder"]


def _params_c(p: PyTwinParams) -> _lib.PyTwinParamsC:
    f = np.float32  # numpy narrows the Python-float constants to float32 against float32 arrays
    return _lib.PyTwinParamsC(f(p.x_min), f(p.x_max), f(p.y_min), f(p.y_max), f(p.z_min), f(p.z_max),
                              f(p.x_div), f(p.y_div), f(p.z_div), p.xn, p.yn, p.zn, int(p.filter_center_car),
                              f(p.ang_res), f(p.vert_res), f(p.velodyne_height),
                              p.front_r_offset, p.front_c_offset, p.front_r_min, p.front_r_max,
                              p.front_c_min, p.front_c_max, p.front_width, p.front_height)


class PyTwinEncoder:
    """One per GPU; batched Python-twin encode of sweeps resident in device memory or in numpy arrays."""

    def __init__(self, device: int = 0):
        self._lib = _lib.load_library()
        self._ctx = ctypes.c_void_p()
        if self._lib.mv3d_ctx_create(int(device), ctypes.byref(self._ctx)) != 0:
            raise RuntimeError("mv3d_ctx_create: " + _lib.last_error())

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

    def encode_device(self, points, offsets, params: PyTwinParams = PYTWIN_KITTI, want_top=True, want_front=True,
                      top_out=None, front_out=None, stream=None):
        """`points`: CUDA float32 `(total, 4)`; returns `(top [B,Xn,Yn,Zn+2] | None, front [B,W,H,3] | None)`."""
        import torch
        assert points.is_cuda and points.dtype == torch.float32 and points.is_contiguous()
        offs = np.ascontiguousarray(offsets, dtype=np.int32)
        batch = len(offs) - 1
        if want_top and top_out is None:
            top_out = torch.empty((batch, *params.top_shape), dtype=torch.float32, device=points.device)
        if want_front and front_out is None:
            front_out = torch.empty((batch, *params.front_shape), dtype=torch.float32, device=points.device)
        if stream is None:
            stream = torch.cuda.current_stream(points.device).cuda_stream
        pc = _params_c(params)
        rc = self._lib.mv3d_pytwin_encode_batch_dev(
            self._ctx, points.data_ptr(), offs.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), batch,
            top_out.data_ptr() if want_top else None, front_out.data_ptr() if want_front else None,
            ctypes.byref(pc), ctypes.c_void_p(stream))
        if rc != 0:
            raise RuntimeError("mv3d_pytwin_encode_batch_dev: " + _lib.last_error())
        return (top_out if want_top else None), (front_out if want_front else None)

    def encode_host(self, points: np.ndarray, offsets, params: PyTwinParams = PYTWIN_KITTI, want_top=True,
                    want_front=True, top_out=None, front_out=None):
        """numpy in, numpy out (H2D, encode, read-back pipelined inside the library). `top_out` / `front_out`:
        C-contiguous float32 arrays with room for the batch, to reuse buffers between calls."""
        points = np.ascontiguousarray(points, dtype=np.float32)
        offs = np.ascontiguousarray(offsets, dtype=np.int32)
        batch = len(offs) - 1

        def buffer(out, shape):
            if out is None:
                return np.empty((batch, *shape), np.float32)
            need = batch * int(np.prod(shape, dtype=np.int64))
            assert out.dtype == np.float32 and out.flags["C_CONTIGUOUS"] and out.size >= need, "output buffer"
            return out.reshape(-1)[:need].reshape((batch, *shape))

        top = buffer(top_out, params.top_shape) if want_top else None
        front = buffer(front_out, params.front_shape) if want_front else None
        pc = _params_c(params)
        rc = self._lib.mv3d_pytwin_encode_batch_host(
            self._ctx, points.ctypes.data, offs.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), batch,
            top.ctypes.data if want_top else None, front.ctypes.data if want_front else None, ctypes.byref(pc))
        if rc != 0:
            raise RuntimeError("mv3d_pytwin_encode_batch_host: " + _lib.last_error())
        return top, front

    def debug_atan2f(self, y, x):
        """Test hook: the library's device restatement of the C library's float atan2f."""
        import torch
        out = torch.empty_like(y)
        rc = self._lib.mv3d_debug_atan2f(self._ctx, y.data_ptr(), x.data_ptr(), out.data_ptr(), y.numel(),
                                         ctypes.c_void_p(torch.cuda.current_stream(y.device).cuda_stream))
        if rc != 0:
            raise RuntimeError("mv3d_debug_atan2f: " + _lib.last_error())
        return out


_default = None


def _encoder() -> PyTwinEncoder:
    global _default
    if _default is None:
        import os
        _default = PyTwinEncoder(int(os.environ.get("MV3D_DEVICE", "0")))
    return _default


def lidar_to_top(lidar, params: PyTwinParams = PYTWIN_KITTI) -> np.ndarray:
    """`src/data.py:296` on the GPU: `(N, 4)` float32 -> float32 `[Xn][Yn][Zn+2]` (flipped, like the reference)."""
    lidar = np.ascontiguousarray(lidar, dtype=np.float32)
    top, _ = _encoder().encode_host(lidar, [0, lidar.shape[0]], params, want_front=False)
    return top[0]


def lidar_to_front_fast(lidar, params: PyTwinParams = PYTWIN_KITTI) -> np.ndarray:
    """`src/data.py:113` on the GPU: `(N, 4)` float32 -> float32 `[FRONT_WIDTH][FRONT_HEIGHT][3]`."""
    lidar = np.ascontiguousarray(lidar, dtype=np.float32)
    _, front = _encoder().encode_host(lidar, [0, lidar.shape[0]], params, want_top=False)
    return front[0]


# The names the reference's loaders import (`from net.utility.front_top_preprocess import lidar_to_top_cuda,
# lidar_to_front_cuda`, src/utils/batch_loading.py:35; call sites :113, :427-428, :661-665). The pycuda twins promise
# what their self-test asserts (src/net/utility/front_top_preprocess.py:213-215, 222-223): every height slice and the
# density channel equal to `lidar_to_top`, the whole front view equal to `lidar_to_front_fast` -- which the functions
# above deliver on every channel, the intensity included.
lidar_to_top_cuda = lidar_to_top
lidar_to_front_cuda = lidar_to_front_fast


class Preprocess:
    """The view-encoding methods of `data.Preprocess` (src/data.py:47-54, 113-180), same names and arguments.
    `lidar_to_top` follows `cfg.USE_CLIDAR_TO_TOP` (src/data.py:49-52): False (the reference's default,
    src/config.py:224) -> the Python encoder's semantics, True -> the C routine through `clidar_to_top`."""

    def __init__(self, params: PyTwinParams = PYTWIN_KITTI, use_clidar_to_top: bool = False):
        self.params = params
        self.use_clidar_to_top = use_clidar_to_top

    def lidar_to_top(self, lidar):
        if self.use_clidar_to_top:
            from .ver3 import clidar_to_top
            return clidar_to_top(lidar)
        return lidar_to_top(lidar, self.params)

    def lidar_to_front_fast(self, lidar):
        return lidar_to_front_fast(lidar, self.params)
