"""Build and load libmv3d_encode.so (the C-ABI library declared in include/mv3d_encode.h).

The library is compiled in-tree with nvcc for sm_100a only. There is no Python or CPU implementation
behind it: if it cannot be built or loaded, importing callers get an exception, not a fallback.
"""
from __future__ import annotations

import ctypes
import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(_HERE, "libmv3d_encode.so")
SOURCES = ["encode_kernels.cu", "sort_strategy.cu", "mv3d_abi.cu", "host_expand.cu", "anchor_kernels.cu", "anchor_abi.cu",
           "sweep_io.cu", "noalloc_front.cu"]
HEADERS = ["encode_kernels.h", "view_math.cuh", "host_expand.h", "ctx.h", "anchor_kernels.h",
           os.path.join("..", "..", "include", "mv3d_encode.h"), os.path.join("..", "..", "include", "mv3d_anchors.h"),
           os.path.join("..", "..", "include", "mv3d_io.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-fmad=false", "-Xcompiler", "-fPIC", "-shared"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libmv3d_encode.so cannot be built (set NVCC=/path/to/nvcc)")


def _stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    built = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS]
    return any(os.path.exists(d) and os.path.getmtime(d) > built for d in deps)


def build_library(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/*.cu into mv3d_b200/libmv3d_encode.so for sm_100a (nvcc cross-compiles without a GPU)."""
    if not force and not _stale():
        return LIB_PATH
    extra = os.environ.get("MV3D_NVCC_EXTRA", "").split()  # tuning experiments, e.g. -DMV3D_POINT_ITEMS=2
    cmd = [_nvcc(), *NVCC_FLAGS, *extra, "-o", LIB_PATH] + [os.path.join(CSRC, s) for s in SOURCES] + ["-lz"]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return LIB_PATH


class TopParamsC(ctypes.Structure):
    _fields_ = [(n, ctypes.c_float) for n in
                ("x_min", "x_max", "y_min", "y_max", "z_min", "z_max", "x_div", "y_div", "z_div")] + \
               [(n, ctypes.c_int) for n in ("xn", "yn", "zn", "flags")]


class FrontParamsC(ctypes.Structure):
    _fields_ = [(n, ctypes.c_float) for n in
                ("phi_min", "phi_max", "theta_min", "theta_max", "phi_div", "theta_div", "c_size", "r_size", "f_size")]


class PyTwinParamsC(ctypes.Structure):
    _fields_ = [(n, ctypes.c_float) for n in
                ("x_min", "x_max", "y_min", "y_max", "z_min", "z_max", "x_div", "y_div", "z_div")] + \
               [(n, ctypes.c_int) for n in ("xn", "yn", "zn", "filter_center_car")] + \
               [(n, ctypes.c_float) for n in ("ang_res", "vert_res", "velodyne_height")] + \
               [(n, ctypes.c_int) for n in ("front_r_offset", "front_c_offset", "front_r_min", "front_r_max",
                                            "front_c_min", "front_c_max", "front_width", "front_height")]


SEM_TOP, SEM_TOPFRONT = 0, 1
STRATEGY_AUTO, STRATEGY_ATOMIC, STRATEGY_TILE, STRATEGY_PATCH = 0, 1, 2, 3
STRATEGY_SORT = 4
STRATEGIES = {"auto": STRATEGY_AUTO, "atomic": STRATEGY_ATOMIC, "tile": STRATEGY_TILE, "patch": STRATEGY_PATCH,
              "sort": STRATEGY_SORT}

# Every symbol include/mv3d_encode.h declares (tests check the built library exports exactly these).
EXPORTED_SYMBOLS = (
    "createTopMaps", "createTopAndFrontMaps", "mv3d_last_error", "mv3d_ctx_create", "mv3d_ctx_destroy",
    "mv3d_encode_batch_dev", "mv3d_encode_batch_host", "mv3d_host_alloc", "mv3d_host_free",
    "mv3d_ctx_launch_count", "mv3d_resolve_strategy", "mv3d_version",
    "mv3d_ctx_set_profiling", "mv3d_kernel_count", "mv3d_kernel_name", "mv3d_ctx_kernel_times",
    "mv3d_front_view_noalloc", "mv3d_debug_host_fill_gbs", "mv3d_debug_host_fill_buffer_gbs", "mv3d_debug_point_keys", "mv3d_debug_expand_split", "mv3d_debug_expand_sorted", "mv3d_debug_selfcheck", "mv3d_debug_selfcheck_trip", "mv3d_ctx_transfer_bytes",
    "mv3d_pytwin_encode_batch_dev", "mv3d_pytwin_encode_batch_host", "mv3d_debug_atan2f", "mv3d_frame_checksums_dev",
    "mv3d_ctx_set_anchors", "mv3d_remove_empty_anchor_dev", "mv3d_remove_empty_anchor_host",
    "mv3d_encode_filter_batch_dev", "mv3d_pytwin_encode_filter_batch_dev",
    "mv3d_io_stat_sweeps", "mv3d_io_read_sweeps", "mv3d_io_pack_sweeps", "mv3d_io_write_npy", "mv3d_io_write_npy_batch", "mv3d_io_npy_info",
    "mv3d_io_read_npy", "mv3d_io_write_npz_compressed", "mv3d_io_write_npz_compressed_batch",
)
ANCHORS_AUTO, ANCHORS_EXACT = 0, 1
ANCHOR_MAX_LINES = 50

_lib = None


def load_library() -> ctypes.CDLL:
    """Load (building first if the sources are newer) and type the C ABI."""
    global _lib
    if _lib is not None:
        return _lib
    if _stale():
        build_library()
    lib = ctypes.CDLL(LIB_PATH)
    cf, ci, vp = ctypes.c_float, ctypes.c_int, ctypes.c_void_p
    lib.createTopMaps.argtypes = [vp, ci, vp] + [cf] * 9 + [ci] * 3
    lib.createTopMaps.restype = None
    lib.createTopAndFrontMaps.argtypes = [vp, ci, vp, vp] + [cf] * 9 + [ci] * 3 + [cf] * 9
    lib.createTopAndFrontMaps.restype = None
    lib.mv3d_last_error.restype = ctypes.c_char_p
    lib.mv3d_version.restype = ctypes.c_char_p
    lib.mv3d_ctx_create.argtypes = [ci, ctypes.POINTER(vp)]
    lib.mv3d_ctx_destroy.argtypes = [vp]
    lib.mv3d_ctx_destroy.restype = None
    lib.mv3d_encode_batch_dev.argtypes = [vp, vp, ctypes.POINTER(ci), ci, vp, vp,
                                          ctypes.POINTER(TopParamsC), ctypes.POINTER(FrontParamsC), ci, ci, vp]
    lib.mv3d_encode_batch_host.argtypes = [vp, vp, ctypes.POINTER(ci), ci, vp, vp,
                                           ctypes.POINTER(TopParamsC), ctypes.POINTER(FrontParamsC), ci, ci]
    lib.mv3d_host_alloc.argtypes = [ctypes.c_size_t]
    lib.mv3d_host_alloc.restype = vp
    lib.mv3d_host_free.argtypes = [vp]
    lib.mv3d_host_free.restype = None
    lib.mv3d_ctx_launch_count.argtypes = [vp]
    lib.mv3d_ctx_launch_count.restype = ctypes.c_longlong
    lib.mv3d_resolve_strategy.argtypes = [vp, ctypes.POINTER(TopParamsC), ci]
    lib.mv3d_ctx_set_profiling.argtypes = [vp, ci]
    lib.mv3d_kernel_name.argtypes = [ci]
    lib.mv3d_kernel_name.restype = ctypes.c_char_p
    lib.mv3d_ctx_kernel_times.argtypes = [vp, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_longlong), ci]
    lib.mv3d_debug_point_keys.argtypes = [vp, vp, ci, ctypes.POINTER(TopParamsC), ctypes.POINTER(FrontParamsC), ci,
                                          vp, vp, vp, vp]
    lib.mv3d_pytwin_encode_batch_dev.argtypes = [vp, vp, ctypes.POINTER(ci), ci, vp, vp, ctypes.POINTER(PyTwinParamsC), vp]
    lib.mv3d_pytwin_encode_batch_host.argtypes = [vp, vp, ctypes.POINTER(ci), ci, vp, vp, ctypes.POINTER(PyTwinParamsC)]
    lib.mv3d_debug_atan2f.argtypes = [vp, vp, vp, vp, ci, vp]
    lib.mv3d_frame_checksums_dev.argtypes = [vp, vp, ci, ctypes.c_longlong, vp, vp]
    lib.mv3d_ctx_transfer_bytes.argtypes = [vp, ctypes.POINTER(ctypes.c_ulonglong), ctypes.POINTER(ctypes.c_ulonglong), ci]
    lib.mv3d_front_view_noalloc.argtypes = [vp, ctypes.POINTER(ci), ci, vp, ctypes.POINTER(TopParamsC),
                                            ctypes.POINTER(FrontParamsC), ci, vp]
    lib.mv3d_debug_host_fill_gbs.argtypes = [ctypes.c_size_t, ci, ci, ci, ci]
    lib.mv3d_debug_host_fill_gbs.restype = ctypes.c_double
    lib.mv3d_debug_host_fill_buffer_gbs.argtypes = [vp, ctypes.c_size_t, ci, ci, ci, ci]
    lib.mv3d_debug_host_fill_buffer_gbs.restype = ctypes.c_double
    lib.mv3d_debug_selfcheck.argtypes = [vp, vp]
    lib.mv3d_debug_selfcheck_trip.argtypes = [vp]
    lib.mv3d_debug_expand_split.argtypes = [vp, ctypes.c_size_t, ci, vp, ctypes.c_uint, vp, ci, ci, ci]
    lib.mv3d_debug_expand_sorted.argtypes = [vp, ctypes.c_size_t, ci, vp, ctypes.c_uint, vp, ci, ci, ci]
    lib.mv3d_ctx_set_anchors.argtypes = [vp, vp, ci]
    lib.mv3d_remove_empty_anchor_dev.argtypes = [vp, vp, ci, ci, ci, ci, cf, ci, vp, vp, vp, vp]
    lib.mv3d_remove_empty_anchor_host.argtypes = [vp, vp, ci, ci, ci, vp, ci, cf, ci, vp, ctypes.POINTER(ci)]
    lib.mv3d_encode_filter_batch_dev.argtypes = [vp, vp, ctypes.POINTER(ci), ci, vp, vp, ctypes.POINTER(TopParamsC),
                                                 ctypes.POINTER(FrontParamsC), ci, ci, cf, ci, vp, vp, vp]
    lib.mv3d_pytwin_encode_filter_batch_dev.argtypes = [vp, vp, ctypes.POINTER(ci), ci, vp, vp,
                                                        ctypes.POINTER(PyTwinParamsC), cf, ci, vp, vp, vp]
    cpp, ll, sz = ctypes.POINTER(ctypes.c_char_p), ctypes.c_longlong, ctypes.c_size_t
    llp = ctypes.POINTER(ll)
    lib.mv3d_io_stat_sweeps.argtypes = [cpp, ci, vp]
    lib.mv3d_io_read_sweeps.argtypes = [cpp, ci, vp, vp, ci]
    lib.mv3d_io_pack_sweeps.argtypes = [vp, ci, vp, vp, ci]
    lib.mv3d_io_write_npy.argtypes = [ctypes.c_char_p, vp, llp, ci]
    lib.mv3d_io_write_npy_batch.argtypes = [cpp, ci, vp, sz, llp, ci, ci]
    lib.mv3d_io_npy_info.argtypes = [ctypes.c_char_p, llp, ctypes.POINTER(ci), llp]
    lib.mv3d_io_read_npy.argtypes = [ctypes.c_char_p, vp, sz]
    lib.mv3d_io_write_npz_compressed.argtypes = [ctypes.c_char_p, ctypes.c_char_p, vp, llp, ci, ci]
    lib.mv3d_io_write_npz_compressed_batch.argtypes = [cpp, ci, ctypes.c_char_p, vp, sz, llp, ci, ci, ci]
    _lib = lib
    return lib


def last_error() -> str:
    return load_library().mv3d_last_error().decode("utf-8", "replace")
