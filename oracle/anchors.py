"""Empty-anchor filter oracle (TEST INFRASTRUCTURE ONLY -- never imported by mv3d_b200).

CPU restatement of `remove_empty_anchor` (src/net/utility/remove_empty_box.py:25-47) and of the kernel it launches
(src/net/utility/remove_empty_box_kernel.cu:3-42). See oracle/anchor_oracle.c for what is pinned and how: the
reference's accumulation across rows is a data race, this is its race-free ascending-row schedule.

`channel_sums` is the C restatement; `channel_sums_numpy` is the same arithmetic in numpy (float32 `cumsum` is a
left-to-right running sum) for small cases; `remove_empty_anchor` finishes the way the reference's host code does,
`np.where(np.sum(index, axis=1) > limit)[0]`.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libanchor_oracle.so")
_SRC = os.path.join(_HERE, "anchor_oracle.c")
REF_CUBIN = os.path.join(_HERE, "_ref", "remove_empty_box_kernel.cubin")
REF_KERNEL = b"_Z12remove_emptyPfPiS_S0_S0_"        # the name remove_empty_box.py:31 looks up
MAX_LINES = 50                                      # grid=(anchors, 50, 1), remove_empty_box.py:43
_lib = None


def build(force: bool = False) -> None:
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(_SRC):
        os.makedirs(os.path.dirname(_SO), exist_ok=True)
        subprocess.run(["gcc", "-O2", "-ffp-contract=off", "-Wall", "-shared", "-fPIC", _SRC, "-o", _SO],
                       check=True, capture_output=True)


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        vp = ctypes.c_void_p
        _lib.anchor_channel_sums.argtypes = [vp, ctypes.c_int, ctypes.c_int, ctypes.c_int, vp, ctypes.c_long, vp]
        _lib.anchor_channel_sums.restype = None
    return _lib


def _check(view, anchors):
    view = np.ascontiguousarray(view, dtype=np.float32)
    anchors = np.ascontiguousarray(anchors, dtype=np.int32).reshape(-1, 4)
    assert view.ndim == 3
    return view, anchors


def channel_sums(view, anchors) -> np.ndarray:
    """`index` of remove_empty_box.py:35-45 after the kernel: float32 [anchors][channels]."""
    view, anchors = _check(view, anchors)
    inds = np.zeros((anchors.shape[0], view.shape[2]), dtype=np.float32)
    lib().anchor_channel_sums(view.ctypes.data, view.shape[0], view.shape[1], view.shape[2], anchors.ctypes.data,
                              anchors.shape[0], inds.ctypes.data)
    return inds


def clamp_box(anchor, xn, yn):
    """(x1, x2, y1, y2) after remove_empty_box_kernel.cu:22-29; x indexes view dimension 0."""
    y1, x1, y2, x2 = (int(v) for v in anchor)
    cl = lambda v, n: min(max(v, 0), n - 1)
    return cl(x1, xn), cl(x2, xn), cl(y1, yn), cl(y2, yn)


def channel_sums_numpy(view, anchors) -> np.ndarray:
    view, anchors = _check(view, anchors)
    xn, yn, c = view.shape
    inds = np.zeros((anchors.shape[0], c), dtype=np.float32)
    for a, box in enumerate(anchors):
        x1, x2, y1, y2 = clamp_box(box, xn, yn)
        lines = min(x2 - x1, MAX_LINES)
        if lines <= 0 or y2 - y1 <= 0:
            continue
        rows = np.cumsum(view[x1:x1 + lines, y1:y2, :], axis=1, dtype=np.float32)[:, -1, :]   # `integration`
        inds[a] = np.cumsum(rows, axis=0, dtype=np.float32)[-1]                                 # `+=` row after row
    return inds


def anchor_totals(view, anchors) -> np.ndarray:
    """`np.sum(index, axis=1)` (remove_empty_box.py:46): numpy's float32 pairwise row sum."""
    return np.sum(channel_sums(view, anchors), axis=1)


def remove_empty_anchor(view, anchors, limit) -> np.ndarray:
    return np.where(anchor_totals(view, anchors) > limit)[0]


def pairwise_sum_f32(a) -> np.float32:
    """numpy's float32 add-reduce of one contiguous row, restated (what the CUDA kernel mirrors):
    0 + pairwise(a), 8 running sums combined as ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)), blocks of <= 128."""
    f32 = np.float32

    def pw(v):
        n = len(v)
        if n < 8:
            r = f32(-0.0)
            for x in v:
                r = f32(r + x)
            return r
        if n <= 128:
            r = [f32(v[j]) for j in range(8)]
            i = 8
            while i < n - (n % 8):
                for j in range(8):
                    r[j] = f32(r[j] + v[i + j])
                i += 8
            res = f32(f32(f32(r[0] + r[1]) + f32(r[2] + r[3])) + f32(f32(r[4] + r[5]) + f32(r[6] + r[7])))
            while i < n:
                res = f32(res + v[i])
                i += 1
            return res
        n2 = n // 2
        n2 -= n2 % 8
        return f32(pw(v[:n2]) + pw(v[n2:]))

    return f32(f32(0) + pw(np.asarray(a, dtype=np.float32)))
