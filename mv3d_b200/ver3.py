"""Host-side mirror of the reference's ver3 ctypes wrappers, bound to libmv3d_encode.so.

`createTopMaps(raw, num, top_flip, top_paras, c_lib_path)` and
`createTopAndFrontMaps(raw, num, top_flip, front_flip, top_paras, front_paras, c_lib_path)` keep the
names, argument order and in-place-output behaviour of the reference wrappers
(`ver3/createTopMaps.py:4-23`, `ver3/createTopAndFrontMaps.py:4-33`), so a script written against those
runs unchanged after `from mv3d_b200.ver3 import *`.

* `raw`        -- `(num, 4)` float32 C-contiguous numpy array (x, y, z, intensity)
* `top_flip`   -- float32 `[Xn][Yn][Zn+2]`, overwritten
* `front_flip` -- float32 `[Rn][Cn][Fn]`, overwritten
* `top_paras`  -- 12 values: 9 floats (x/y/z MIN, MAX, then the three DIVISIONs) and `Xn, Yn, Zn`
* `front_paras`-- 9 values: phi MIN/MAX, theta MIN/MAX, phi/theta DIVISION, then `Rn, Cn, Fn`; the C symbol
                  receives the sizes as `C_SIZE=Cn, R_SIZE=Rn, F_SIZE=Fn` exactly like the reference (`:32`)
* `c_lib_path` -- optional; defaults to this package's library (the reference made it mandatory)

The caller-side flip stays with the caller: `np.flipud(np.fliplr(top_flip))` (`src/data.py:290`).
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import _lib

__all__ = ["createTopMaps", "createTopAndFrontMaps", "clidar_to_top"]

_F, _I, _P = ctypes.c_float, ctypes.c_int, ctypes.c_void_p


def _resolve(c_lib_path):
    if c_lib_path in (None, _lib.LIB_PATH):
        return _lib.load_library()
    foreign = ctypes.cdll.LoadLibrary(c_lib_path)  # e.g. the reference's own .so, for side-by-side runs
    foreign.createTopMaps.argtypes = [_P, _I, _P] + [_F] * 9 + [_I] * 3
    if hasattr(foreign, "createTopAndFrontMaps"):
        foreign.createTopAndFrontMaps.argtypes = [_P, _I, _P, _P] + [_F] * 9 + [_I] * 3 + [_F] * 9
    return foreign


def _check(arr, name, min_floats):
    if not isinstance(arr, np.ndarray) or arr.dtype != np.float32 or not arr.flags["C_CONTIGUOUS"]:
        raise TypeError(f"{name} must be a C-contiguous float32 numpy array")
    if arr.size < min_floats:
        raise ValueError(f"{name} holds {arr.size} floats, the call needs {min_floats}")


def createTopMaps(raw, num, top_flip, top_paras, c_lib_path=None):
    """Bird's-eye-view encode of one sweep through the `createTopMaps` C symbol."""
    bounds, grid = tuple(top_paras[:9]), tuple(int(v) for v in top_paras[9:12])
    _check(raw, "raw", 4 * num)
    _check(top_flip, "top_flip", grid[0] * grid[1] * (grid[2] + 2))
    lib = _resolve(c_lib_path)
    lib.createTopMaps(raw.ctypes.data, int(num), top_flip.ctypes.data, *map(float, bounds), *grid)
    err = _lib.last_error() if lib is _lib.load_library() else ""
    if err:
        raise RuntimeError("createTopMaps: " + err)


def createTopAndFrontMaps(raw, num, top_flip, front_flip, top_paras, front_paras, c_lib_path=None):
    """Bird's-eye-view + front-view encode of one sweep through the `createTopAndFrontMaps` C symbol."""
    bounds, grid = tuple(top_paras[:9]), tuple(int(v) for v in top_paras[9:12])
    angles = tuple(front_paras[:6])
    rn, cn, fn = (int(v) for v in front_paras[6:9])
    _check(raw, "raw", 4 * num)
    _check(top_flip, "top_flip", grid[0] * grid[1] * (grid[2] + 2))
    _check(front_flip, "front_flip", rn * cn * fn)
    lib = _resolve(c_lib_path)
    lib.createTopAndFrontMaps(raw.ctypes.data, int(num), top_flip.ctypes.data, front_flip.ctypes.data,
                              *map(float, bounds), *grid, *map(float, angles), float(cn), float(rn), float(fn))
    err = _lib.last_error() if lib is _lib.load_library() else ""
    if err:
        raise RuntimeError("createTopAndFrontMaps: " + err)


_clidar_encoder = None


def clidar_to_top(lidar, top=None):
    """`src/data.py:265-291` on the new library: [centre-car filter,] `createTopMaps`, `np.flipud(np.fliplr(.))`.

    `top`: a `TopParams`; default `CLIDAR_KITTI` (the reference picks the TOP_* constants and the filter from
    `cfg.DATA_SETS_TYPE`: see `params.CLIDAR_PRESETS`). With `top.flip` / `top.filter_center_car` set, both steps
    happen inside the encode (MV3D_TOP_FLIP_XY / MV3D_TOP_FILTER_CENTER_CAR) and the result is a contiguous array;
    a parameter set without the flags is encoded as the C routine does and flipped as a numpy view, like the
    reference.

    Deviation from the reference, on purpose: the output is float32, the type the C symbol writes;
    the reference allocates float64 there and reads back garbage.
    """
    global _clidar_encoder
    from .params import CLIDAR_KITTI
    top = top or CLIDAR_KITTI
    lidar = np.ascontiguousarray(lidar, dtype=np.float32)
    if top.flags:
        if _clidar_encoder is None:
            import os
            from .encoder import ViewEncoder
            _clidar_encoder = ViewEncoder(int(os.environ.get("MV3D_DEVICE", "0")), semantics="top")
        out, _ = _clidar_encoder.encode_host(lidar, [0, lidar.shape[0]], top)
        return out[0]
    top_flip = np.ones(top.shape, dtype=np.float32)
    createTopMaps(lidar, lidar.shape[0], top_flip, top.as_top_paras())
    return np.flipud(np.fliplr(top_flip))
