"""Python-twin oracle (TEST INFRASTRUCTURE ONLY -- never imported by mv3d_b200).

numpy restatement of the reference's PYTHON view encoders, the ones `src/preprocess_for_mv3d.py:30-32` and the
training loaders actually run:

* `lidar_to_top`         src/data.py:296-367   -> float32 [Xn][Yn][Zn+2], written already flipped
* `lidar_to_front_fast`  src/data.py:113-180   -> float32 [FRONT_WIDTH][FRONT_HEIGHT][3], mean per pixel

Parity status: PINNED by execution. tests/test_pytwin_oracle.py runs the reference's own functions (compiled out
of /root/reference/src/data.py at test time by tests/golden/pytwin_reference.py) against this restatement, and
tests/golden/pytwin_golden.npz holds outputs of that run for machines without /root/reference.

float32 `np.arctan2` is taken from the C library's `atan2f` (oracle/pytwin_oracle.c), as the numpy of the
reference's time did; a current numpy with AVX-512 dispatch returns different last bits (see that file's header).
"""
from __future__ import annotations

import ctypes
import math
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libpytwin_oracle.so")
_SRC = os.path.join(_HERE, "pytwin_oracle.c")
f32 = np.float32
_lib = None


def build(force: bool = False) -> None:
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
        os.makedirs(os.path.dirname(_SO), exist_ok=True)
        subprocess.run(["gcc", "-O2", "-ffp-contract=off", "-Wall", "-shared", "-fPIC", _SRC, "-o", _SO, "-lm"],
                       check=True, capture_output=True)


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        vp = ctypes.c_void_p
        for name in ("fdlibm_atan2f_array", "libm_atan2f_array"):
            getattr(_lib, name).argtypes = [vp, vp, vp, ctypes.c_long]
        _lib.npy_floor_dividef_array.argtypes = [vp, ctypes.c_float, vp, ctypes.c_long]
    return _lib


def atan2f(y, x, impl: str = "libm") -> np.ndarray:
    y = np.ascontiguousarray(y, dtype=np.float32)
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty_like(y)
    getattr(lib(), impl + "_atan2f_array")(y.ctypes.data, x.ctypes.data, out.ctypes.data, y.size)
    return out


def floor_divide_f32(a, b) -> np.ndarray:
    """The C restatement of numpy's float32 `//` (what the CUDA port mirrors); equals `a // np.float32(b)`."""
    a = np.ascontiguousarray(a, dtype=np.float32)
    out = np.empty_like(a)
    lib().npy_floor_dividef_array(a.ctypes.data, ctypes.c_float(b), out.ctypes.data, a.size)
    return out


def crop(lidar, p) -> np.ndarray:
    """Open-interval crop on all three axes, in float32 (src/data.py:298-311 = :130-143). Returns a boolean mask."""
    x, y, z = lidar[:, 0], lidar[:, 1], lidar[:, 2]
    return ((x > f32(p.x_min)) & (x < f32(p.x_max)) & (y > f32(p.y_min)) & (y < f32(p.y_max)) &
            (z > f32(p.z_min)) & (z < f32(p.z_max)))


def center_car(lidar) -> np.ndarray:
    """Mask of `filter_center_car` (src/data.py:224-227): keep |x| > 4.7/2 or |y| > 2.1/2."""
    return (np.abs(lidar[:, 0]) > f32(4.7 / 2)) | (np.abs(lidar[:, 1]) > f32(2.1 / 2))


def density_table() -> np.ndarray:
    """`min(1, np.log(count+1)/math.log(32))` (src/data.py:354) in double, stored as float32; index = count."""
    return np.array([min(1.0, math.log(c + 1) / math.log(32)) for c in range(32)], dtype=np.float32)


def lidar_to_top(lidar, p) -> np.ndarray:
    lidar = np.ascontiguousarray(lidar, dtype=np.float32)
    keep = crop(lidar, p)
    if p.filter_center_car:                                               # :313-314
        keep &= center_car(lidar)
    pts = lidar[keep]                                                     # input order is preserved
    xn, yn, zn = p.xn, p.yn, p.zn
    top = np.zeros((xn, yn, zn + 2), dtype=np.float32)
    if len(pts) == 0:
        return top
    qx = ((pts[:, 0] - f32(p.x_min)) // f32(p.x_div)).astype(np.int32)    # :321 numpy float32 floor_divide
    qy = ((pts[:, 1] - f32(p.y_min)) // f32(p.y_div)).astype(np.int32)    # :322
    qz = (pts[:, 2] - f32(p.z_min)) / f32(p.z_div)                        # :324 float32, kept fractional
    inside = (qx >= 0) & (qx < xn) & (qy >= 0) & (qy < yn)                # the loops :341,:347 only visit these
    qx, qy, qz, inten = qx[inside], qy[inside], qz[inside], pts[inside, 3]
    cell = (xn - 1 - qx).astype(np.int64) * yn + (yn - 1 - qy)            # written at top[-x-1, -y-1] :345,:352
    ncell = xn * yn
    flat = top.reshape(ncell, zn + 2)
    # density :354
    count = np.bincount(cell, minlength=ncell)
    lut = density_table()
    touched = np.flatnonzero(count)
    flat[touched, zn + 1] = lut[np.minimum(count[touched], 31)]
    # intensity of the first point (crop order) with the largest qz :355-356 (np.argmax returns the first maximum)
    order = np.lexsort((np.arange(len(cell)), -qz.astype(np.float64), cell))
    first = np.ones(len(order), dtype=bool)
    first[1:] = cell[order][1:] != cell[order][:-1]
    win = order[first]
    flat[cell[win], zn] = inten[win]
    # height slices :358-366: slice z takes every point with z <= qz <= z+1 (both ends inclusive), value max(qz) - z
    qzd = qz.astype(np.float64)
    raw = np.full((ncell, zn), -1.0)
    s = np.floor(qzd).astype(np.int64)
    own = (s >= 0) & (s < zn)
    np.maximum.at(raw, (cell[own], s[own]), qzd[own])
    lower = (qzd == s) & (s >= 1) & (s - 1 < zn)                          # an integral qz also closes the slice below
    np.maximum.at(raw, (cell[lower], s[lower] - 1), qzd[lower])
    ci, zi = np.nonzero(raw >= 0)
    flat[ci, zi] = np.maximum(0.0, raw[ci, zi] - zi).astype(np.float32)
    return top


def front_bins(pts, p):
    """(c, r) of `to_front` (src/data.py:121-126): float32 atan2 / float32 resolution, truncated to int32."""
    x, y, z = pts[:, 0], pts[:, 1], pts[:, 2]
    c = (atan2f(y, x) / f32(p.ang_res)).astype(np.int32)
    r = (atan2f(z, np.sqrt(x ** 2 + y ** 2)) / f32(p.vert_res)).astype(np.int32)
    return c, r


def lidar_to_front_fast(lidar, p) -> np.ndarray:
    lidar = np.ascontiguousarray(lidar, dtype=np.float32)
    pts = lidar[crop(lidar, p)]                                           # :130-143 (no centre-car filter here)
    front = np.zeros((p.front_width, p.front_height, 3), dtype=np.float32)
    if len(pts) == 0:
        return front
    c, r = front_bins(pts, p)
    ok = (p.front_c_min < c) & (c < p.front_c_max) & (p.front_r_min < r) & (r < p.front_r_max)   # :146-153
    c = c + p.front_c_offset                                              # :155-156
    r = r + p.front_r_offset
    ok &= (0 <= c) & (c < p.front_width) & (0 <= r) & (r < p.front_height)                       # :159-166
    pts, c, r = pts[ok], c[ok], r[ok]
    height = np.clip(pts[:, 2] + f32(p.velodyne_height), 0, None).astype(np.float32)             # :115-116
    sq = pts ** 2
    dist = np.sqrt(((sq[:, 0] + sq[:, 1]) + sq[:, 2]) + sq[:, 3])          # :117-118 (intensity is inside the norm, sic)
    vals = np.stack([height, dist.astype(np.float32), pts[:, 3]], axis=1)
    pix = c.astype(np.int64) * p.front_height + r
    flat = front.reshape(-1, 3)
    # `front[c, r, :] += x[2:]` row by row (:172-177): a float32 running sum in input order, then / count (:178)
    acc = np.zeros_like(flat)
    cnt = np.zeros(flat.shape[0], dtype=np.float32)
    for i in np.argsort(pix, kind="stable"):                              # per pixel, input order
        acc[pix[i]] = acc[pix[i]] + vals[i]
        cnt[pix[i]] += 1
    hit = cnt > 0
    flat[hit] = acc[hit] / cnt[hit, None]
    return front
