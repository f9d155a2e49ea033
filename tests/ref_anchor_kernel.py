"""Launches the REFERENCE's own empty-anchor kernel on the GPU box (test infrastructure).

oracle/Makefile compiles src/net/utility/remove_empty_box_kernel.cu, unmodified and from where it lies, into
oracle/_ref/remove_empty_box_kernel.cubin (sm_100a instead of the shipped sm_61 cubin); this module loads that cubin
by the mangled name the reference looks up and launches it exactly as remove_empty_box.py:36-45 does --
grid=(anchors, 50, 1), block=(channels, 1, 1), the five pointer arguments -- through the CUDA driver API
(cuda-python instead of pycuda, which is not in this image).
"""
import ctypes
import os

import numpy as np

from oracle.anchors import MAX_LINES, REF_CUBIN, REF_KERNEL


def available() -> bool:
    if not os.path.exists(REF_CUBIN):
        return False
    try:
        from cuda.bindings import driver  # noqa: F401
    except Exception:
        return False
    return True


def _ok(res):
    err = res[0]
    if int(err) != 0:
        raise RuntimeError(f"CUDA driver error {err}")
    return res[1] if len(res) == 2 else res[1:]


_func = None


def _function():
    global _func
    if _func is None:
        from cuda.bindings import driver
        module = _ok(driver.cuModuleLoadData(open(REF_CUBIN, "rb").read()))
        _func = (_ok(driver.cuModuleGetFunction(module, REF_KERNEL)), module)
    return _func[0]


def channel_sums(view: np.ndarray, anchors: np.ndarray) -> np.ndarray:
    """`index` after the reference kernel ran: float32 [anchors][channels] (racy across rows, like the reference)."""
    import torch
    from cuda.bindings import driver
    torch.cuda.init()
    torch.zeros(1, device="cuda")  # makes torch's primary context current for the driver calls below
    view = np.ascontiguousarray(view, dtype=np.float32)
    anchors = np.ascontiguousarray(anchors, dtype=np.int32).reshape(-1, 4)
    n, c = anchors.shape[0], view.shape[2]
    assert 0 < c <= 1024 and n > 0
    d_view = torch.from_numpy(view).cuda()
    d_anchors = torch.from_numpy(anchors).cuda()
    d_inds = torch.zeros((n, c), dtype=torch.float32, device="cuda")
    d_ashape = torch.tensor(anchors.shape, dtype=torch.int32, device="cuda")
    d_vshape = torch.tensor(view.shape, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ptrs = tuple(int(t.data_ptr()) for t in (d_inds, d_anchors, d_view, d_ashape, d_vshape))
    args = (ptrs, (ctypes.c_void_p,) * 5)
    _ok(driver.cuLaunchKernel(_function(), n, MAX_LINES, 1, c, 1, 1, 0, 0, args, 0))
    _ok(driver.cuCtxSynchronize())
    return d_inds.cpu().numpy()
