"""Parity oracle loaders (TEST INFRASTRUCTURE ONLY -- never imported by mv3d_b200).

Two CPU implementations of the view encoder, both behind ctypes:

* `restated`  -- oracle/view_oracle.c, our flat-array restatement (oracle/_build/libview_oracle.so);
* `reference` -- the reference's own ver3 C sources compiled unmodified (oracle/_ref/*.so), when present.

Both are called the way `ver3/createTopMaps.py:15-23` / `ver3/createTopAndFrontMaps.py:20-33` call the
reference. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use this.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
RESTATED_SO = os.path.join(_HERE, "_build", "libview_oracle.so")
PROBE_SO = os.path.join(_HERE, "_build", "libref_probe.so")
REF_TOP_SO = os.path.join(_HERE, "_ref", "LidarTopPreprocess.so")
REF_TOPFRONT_SO = os.path.join(_HERE, "_ref", "LidarTopAndFrontPreprocess.so")

_cf, _ci, _vp = ctypes.c_float, ctypes.c_int, ctypes.c_void_p
_TOP_ARGTYPES = [_vp, _ci, _vp] + [_cf] * 9 + [_ci] * 3
_TOPFRONT_ARGTYPES = [_vp, _ci, _vp, _vp] + [_cf] * 9 + [_ci] * 3 + [_cf] * 9


def build(force: bool = False) -> None:
    """Compile the restatement (always) and oracle/_ref (only where /root/reference exists)."""
    if force or not os.path.exists(RESTATED_SO) or \
            os.path.getmtime(RESTATED_SO) < os.path.getmtime(os.path.join(_HERE, "view_oracle.c")):
        subprocess.run(["make", "-C", _HERE, "_build/libview_oracle.so"], check=True, capture_output=True)
    if force or not os.path.exists(PROBE_SO) or \
            os.path.getmtime(PROBE_SO) < os.path.getmtime(os.path.join(_HERE, "ref_probe.c")):
        subprocess.run(["make", "-C", _HERE, "_build/libref_probe.so"], check=True, capture_output=True)
    if force or not reference_available():
        subprocess.run(["make", "-C", _HERE, "ref"], check=True, capture_output=True)


def reference_available() -> bool:
    return os.path.exists(REF_TOP_SO) and os.path.exists(REF_TOPFRONT_SO)


_libs = {}


def _lib(path):
    if path not in _libs:
        _libs[path] = ctypes.CDLL(path)
    return _libs[path]


def _points(raw):
    raw = np.ascontiguousarray(raw, dtype=np.float32)
    assert raw.ndim == 2 and raw.shape[1] == 4, "points must be (N, 4) float32"
    return raw


def _front_alloc(front):
    # The reference stores one row + one pixel past the end when r_coord == R_SIZE:
    # give it room, return only the in-bounds image.
    n = front.rn * front.cn * front.fn
    buf = np.ones(n + front.cn * front.fn + 2 * front.fn + 16, dtype=np.float32)
    return buf, n


def top_maps(raw, top, impl: str = "restated") -> np.ndarray:
    """`createTopMaps` semantics -> float32 `[Xn][Yn][Zn+2]` (un-flipped, as the ABI leaves it)."""
    raw = _points(raw)
    out = np.ones(top.shape, dtype=np.float32)  # callee overwrites everything
    if impl == "restated":
        fn = _lib(RESTATED_SO).oracle_createTopMaps
    elif impl == "reference":
        fn = _lib(REF_TOP_SO).createTopMaps
    else:
        raise ValueError(impl)
    fn.argtypes = _TOP_ARGTYPES
    fn(raw.ctypes.data, raw.shape[0], out.ctypes.data, *top.floats, top.xn, top.yn, top.zn)
    return out


def top_and_front_maps(raw, top, front, impl: str = "restated"):
    """`createTopAndFrontMaps` semantics -> (top `[Xn][Yn][Zn+2]`, front `[R][C][F]`), float32."""
    raw = _points(raw)
    out = np.ones(top.shape, dtype=np.float32)
    fbuf, n = _front_alloc(front)
    if impl == "restated":
        fn = _lib(RESTATED_SO).oracle_createTopAndFrontMaps
    elif impl == "reference":
        fn = _lib(REF_TOPFRONT_SO).createTopAndFrontMaps
    else:
        raise ValueError(impl)
    fn.argtypes = _TOPFRONT_ARGTYPES
    fn(raw.ctypes.data, raw.shape[0], out.ctypes.data, fbuf.ctypes.data, *top.floats,
       top.xn, top.yn, top.zn, front.phi_min, front.phi_max, front.theta_min, front.theta_max,
       front.phi_div, front.theta_div, float(front.cn), float(front.rn), float(front.fn))
    return out, fbuf[:n].reshape(front.shape).copy()


def point_keys(raw, top, front=None, semantics: str = "topfront"):
    """Per-point (cell, front pixel) decisions of the restatement; -1 = rejected / not stored."""
    raw = _points(raw)
    n = raw.shape[0]
    cell = np.empty(n, dtype=np.int32)
    pix = np.empty(n, dtype=np.int32) if front is not None else None
    fn = _lib(RESTATED_SO).oracle_point_keys
    fn.argtypes = [_vp, _ci, _ci] + [_cf] * 6 + [_ci] * 3 + [_cf] * 4 + [_vp, _vp]
    fn(raw.ctypes.data, n, {"top": 0, "topfront": 1}[semantics], top.x_min, top.y_min, top.z_min,
       top.x_div, top.y_div, top.z_div, top.xn, top.yn, top.zn,
       front.phi_div if front else 1.0, front.theta_div if front else 1.0,
       float(front.cn) if front else 1.0, float(front.rn) if front else 1.0,
       cell.ctypes.data, pix.ctypes.data if pix is not None else None)
    return cell, pix


def reference_point_keys(raw, top, front):
    """Per-point (cell, front pixel) decisions of the REFERENCE build, one point per call (oracle/ref_probe.c).
    -1 = gated out / nothing stored inside the image. Needs oracle/_ref; use a small bird's-eye-view grid, the
    reference allocates it afresh on every call."""
    raw = _points(raw)
    n = raw.shape[0]
    guess = point_keys(raw, top, front)[1]
    cell = np.empty(n, dtype=np.int32)
    pix = np.empty(n, dtype=np.int32)
    fn = _lib(PROBE_SO).ref_probe_points
    fn.argtypes = [ctypes.c_char_p, _vp, _ci] + [_cf] * 9 + [_ci] * 3 + [_cf] * 2 + [_ci] * 2 + [_vp] * 3
    fn.restype = _ci
    rc = fn(REF_TOPFRONT_SO.encode(), raw.ctypes.data, n, *top.floats, top.xn, top.yn, top.zn, front.phi_div,
            front.theta_div, front.cn, front.rn, guess.ctypes.data, cell.ctypes.data, pix.ctypes.data)
    if rc:
        raise RuntimeError("ref_probe_points: cannot load " + REF_TOPFRONT_SO)
    return cell, pix


def density_table() -> np.ndarray:
    out = np.zeros(32, dtype=np.float32)
    fn = _lib(RESTATED_SO).oracle_density_table
    fn.argtypes = [_vp]
    fn(out.ctypes.data)
    return out
