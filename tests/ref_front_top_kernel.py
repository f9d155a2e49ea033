"""Launches the REFERENCE's own GPU view encoder on the GPU box (test / bench infrastructure, never product).

oracle/Makefile compiles src/net/utility/front_top_kernel.cu, unmodified and from where it lies, into
oracle/_ref/front_top_kernel.cubin (sm_100a instead of the shipped sm_61 cubin). This module loads the four kernels
by the mangled names the reference looks up (front_top_preprocess.py:33-34,103-104) and drives them with the
reference's own launch shapes and host steps (front_top_preprocess.py:29-193): crop, voxelise, np.lexsort, upload,
`lidar_to_top` with grid=(N,1,1) block=(channels,1,1), the single-thread `lidar_to_top_density`, the single-thread
`lidar_to_front_add_points`, the three-thread `lidar_to_front_fill_front`, download, the numpy epilogues.
cuda-python's driver API stands in for pycuda (not in this image); pycuda's In/InOut are synchronous copies around a
synchronous launch, which is what the explicit copies + cuCtxSynchronize below amount to.

float32 np.arctan2 comes from oracle.pytwin.atan2f (the C library's atan2f, what the numpy of the reference's time
called); everything else is the numpy the reference wrote.
"""
import ctypes
import math
import os
import time

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_CUBIN = os.path.join(os.path.dirname(_HERE), "oracle", "_ref", "front_top_kernel.cubin")
K_TOP = b"_Z12lidar_to_topPfPiS0_S0_S_S_S0_"
K_DENSITY = b"_Z20lidar_to_top_densityPfPiS0_S0_S0_"
K_ADD = b"_Z25lidar_to_front_add_pointsPiS_S_S_"
K_FILL = b"_Z25lidar_to_front_fill_frontPfS_PiS0_"


def available() -> bool:
    if not os.path.exists(REF_CUBIN):
        return False
    try:
        from cuda.bindings import driver  # noqa: F401
    except Exception:
        return False
    return True


def _ok(res):
    if int(res[0]) != 0:
        raise RuntimeError(f"CUDA driver error {res[0]}")
    return res[1] if len(res) == 2 else res[1:]


_funcs = None


def _functions():
    global _funcs
    if _funcs is None:
        import torch
        from cuda.bindings import driver
        torch.cuda.init()
        torch.zeros(1, device="cuda")  # torch's primary context becomes current for the driver calls
        module = _ok(driver.cuModuleLoadData(open(REF_CUBIN, "rb").read()))
        _funcs = {k: _ok(driver.cuModuleGetFunction(module, k)) for k in (K_TOP, K_DENSITY, K_ADD, K_FILL)}
        _funcs["module"] = module
    return _funcs


def _launch(name, grid, block, tensors):
    from cuda.bindings import driver
    ptrs = tuple(int(t.data_ptr()) for t in tensors)
    _ok(driver.cuLaunchKernel(_functions()[name], grid, 1, 1, block, 1, 1, 0, 0, (ptrs, (ctypes.c_void_p,) * len(ptrs)), 0))
    _ok(driver.cuCtxSynchronize())


def _crop(lidar, p):
    """front_top_preprocess.py:36-50 / :121-134: six successive open-interval filters."""
    for axis, lo, hi in ((0, p.x_min, p.x_max), (1, p.y_min, p.y_max), (2, p.z_min, p.z_max)):
        lidar = lidar[np.where(lidar[:, axis] > lo)]
        lidar = lidar[np.where(lidar[:, axis] < hi)]
    return lidar


def lidar_to_top_cuda(lidar: np.ndarray, p, timing: dict | None = None) -> np.ndarray:
    """front_top_preprocess.py:29-97 with the reference kernels; `p` is a PyTwinParams (the TOP_* constants)."""
    import torch
    t0 = time.perf_counter()
    lidar = _crop(np.copy(lidar), p)
    height, width, channel = p.xn, p.yn, p.zn + 2                                              # :52-58
    top = np.zeros((height, width, channel), np.float32)
    top_density = np.zeros((height, width, 1), np.float32)
    top_shape = np.array(top.shape).astype(np.int32)
    lidar_shape = np.array(lidar.shape).astype(np.int32)
    lidar[:, 0] = ((lidar[:, 0] - p.x_min) // p.x_div).astype(np.int32)                        # :65-67
    lidar[:, 1] = ((lidar[:, 1] - p.y_min) // p.y_div).astype(np.int32)
    lidar[:, 2] = (lidar[:, 2] - p.z_min) / p.z_div
    lidar = lidar[np.lexsort((lidar[:, 2], lidar[:, 1], lidar[:, 0])), :]                      # :69
    lidar_x = np.ascontiguousarray(lidar[:, 0].astype(np.int32))
    lidar_y = np.ascontiguousarray(lidar[:, 1].astype(np.int32))
    lidar_z = np.ascontiguousarray(lidar[:, 2])
    lidar_i = np.ascontiguousarray(lidar[:, 3])
    t1 = time.perf_counter()
    n = int(lidar_shape[0])
    if n > 0:
        d = [torch.from_numpy(a).cuda() for a in (top, top_shape, lidar_x, lidar_y, lidar_z, lidar_i, lidar_shape)]
        _launch(K_TOP, n, channel, d)                                                           # :75-86
        top = d[0].cpu().numpy()
        dd = torch.from_numpy(top_density).cuda()
        _launch(K_DENSITY, 1, 1, [dd, d[2], d[3], d[6], d[1]])                                 # :87-95
        top_density = dd.cpu().numpy()
    t2 = time.perf_counter()
    top_density = (np.log(top_density.astype(np.int32) + 1) / math.log(32)).clip(max=1).astype(np.float32)  # :96
    out = np.dstack([top[:, :, :-1], top_density])
    if timing is not None:
        timing["top_host_s"] = timing.get("top_host_s", 0.0) + (t1 - t0) + (time.perf_counter() - t2)
        timing["top_gpu_s"] = timing.get("top_gpu_s", 0.0) + (t2 - t1)
    return out


def lidar_to_front_cuda(lidar: np.ndarray, p, timing: dict | None = None) -> np.ndarray:
    """front_top_preprocess.py:99-193 with the reference kernels."""
    import torch
    from oracle.pytwin import atan2f
    t0 = time.perf_counter()
    f32 = np.float32
    lidar = _crop(lidar, p)
    # to_front (:113-118): float32 arctan2 over a double resolution, truncated by the int32 array constructor
    c = atan2f(lidar[:, 1], lidar[:, 0]) / p.ang_res
    r = atan2f(lidar[:, 2], np.sqrt(lidar[:, 0] ** 2 + lidar[:, 1] ** 2)) / p.vert_res
    points = np.array([c, r], dtype=np.int32).T
    for col, lo, hi in ((0, p.front_c_min, p.front_c_max), (1, p.front_r_min, p.front_r_max)):   # :137-144
        ind = np.where(lo < points[:, col]); points, lidar = points[ind], lidar[ind]
        ind = np.where(points[:, col] < hi); points, lidar = points[ind], lidar[ind]
    points[:, 0] += int(p.front_c_offset)
    points[:, 1] += int(p.front_r_offset)
    for col, hi in ((0, p.front_width), (1, p.front_height)):                                   # :150-157
        ind = np.where(0 <= points[:, col]); points, lidar = points[ind], lidar[ind]
        ind = np.where(points[:, col] < hi); points, lidar = points[ind], lidar[ind]
    idx = np.lexsort((points[:, 1], points[:, 0]))                                              # :160-162
    points, lidar = points[idx, :], lidar[idx, :]
    front = np.zeros((p.front_width, p.front_height, 3), np.float32)
    weight_mask = np.zeros_like(front[:, :, 0]).astype(np.int32)
    height = np.clip(lidar[:, 2] + p.velodyne_height, a_min=0, a_max=None).astype(f32).reshape((-1, 1))
    distance = np.sqrt(np.sum(lidar ** 2, axis=1)).astype(f32).reshape((-1, 1))
    intensity = lidar[:, 3].astype(f32).reshape((-1, 1))
    buf = np.hstack((points, height, distance, intensity)).astype(f32)                          # :171
    points = np.ascontiguousarray(points)
    t1 = time.perf_counter()
    if points.shape[0] > 0:
        d_mask = torch.from_numpy(weight_mask).cuda()
        d_points = torch.from_numpy(points).cuda()
        _launch(K_ADD, 1, 1, [d_mask, d_points, torch.tensor(weight_mask.shape, dtype=torch.int32).cuda(),
                              torch.tensor(points.shape, dtype=torch.int32).cuda()])            # :174-181
        weight_mask = d_mask.cpu().numpy()
    weight_mask[weight_mask == 0] = 1
    if points.shape[0] > 0:
        d_front = torch.from_numpy(front).cuda()
        _launch(K_FILL, 1, 3, [d_front, torch.from_numpy(buf).cuda(), torch.tensor(front.shape, dtype=torch.int32).cuda(),
                               torch.tensor(buf.shape, dtype=torch.int32).cuda()])              # :183-190
        front = d_front.cpu().numpy()
    t2 = time.perf_counter()
    front /= weight_mask[:, :, np.newaxis]
    if timing is not None:
        timing["front_host_s"] = timing.get("front_host_s", 0.0) + (t1 - t0) + (time.perf_counter() - t2)
        timing["front_gpu_s"] = timing.get("front_gpu_s", 0.0) + (t2 - t1)
    return front
