"""Runs the REFERENCE's own Python encoders, from where they lie under /root/reference (build container only).

`src/data.py` cannot be imported whole here (it pulls in pykitti, OpenCV drawing helpers, the detector's box code), so
this module compiles just the three functions on the path out of the reference file at run time --
`lidar_to_top` (src/data.py:296-367), `filter_center_car` (:224-227) and `Preprocess.lidar_to_front_fast` (:113-180)
-- and executes them with the module-level names they expect (`np`, `numpy`, `math`, `cfg`, `TOP_*`), filled from a
`PyTwinParams`. No reference source is copied into this repository.

numpy >= 1.22 answers float32 `np.arctan2` from an AVX-512 loop; the numpy of the reference's time (1.13) called the
C library's `atan2f`. Set NPY_DISABLE_CPU_FEATURES (see `LIBM_ENV`) before importing numpy to get the latter; the
committed fixtures were generated that way (`make_pytwin_golden.py` re-executes itself with it).
"""
import ast
import math
import os
import types

REFERENCE_DATA_PY = "/root/reference/src/data.py"
LIBM_ENV = {"NPY_DISABLE_CPU_FEATURES": "AVX512F AVX512CD AVX512_SKX AVX512_CLX AVX512_CNL AVX512_ICL AVX512_SPR AVX2 FMA3"}


def available() -> bool:
    return os.path.exists(REFERENCE_DATA_PY)


def _extract(tree, name, inside_class=None):
    body = tree.body
    if inside_class:
        body = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == inside_class).body
    return next(n for n in body if isinstance(n, ast.FunctionDef) and n.name == name)


class _NumpyCompat:
    """numpy as the reference's code expects it: `np.object` (removed in numpy 1.24, used at src/data.py:402)."""
    object = object

    def __getattr__(self, name):
        import numpy
        return getattr(numpy, name)


def load(params, others: bool = False):
    """Returns (lidar_to_top, lidar_to_front_fast) of the reference, bound to `params`; with `others=True` the two
    remaining twins instead: (lidar_to_top_fast, src/data.py:369-444; Preprocess.lidar_to_front, :56-111)."""
    import numpy as np
    src = open(REFERENCE_DATA_PY).read()
    tree = ast.parse(src)
    if others:
        funcs = [_extract(tree, "filter_center_car"), _extract(tree, "lidar_to_top_fast"),
                 _extract(tree, "lidar_to_front", inside_class="Preprocess")]
    else:
        funcs = [_extract(tree, "filter_center_car"), _extract(tree, "lidar_to_top"),
                 _extract(tree, "lidar_to_front_fast", inside_class="Preprocess")]
    module = ast.Module(body=funcs, type_ignores=[])
    cfg = types.SimpleNamespace(
        DATA_SETS_TYPE="didi2" if params.filter_center_car else "kitti",
        VELODYNE_ANGULAR_RESOLUTION=params.ang_res, VELODYNE_VERTICAL_RESOLUTION=params.vert_res,
        VELODYNE_HEIGHT=params.velodyne_height, FRONT_R_OFFSET=params.front_r_offset,
        FRONT_C_OFFSET=params.front_c_offset, FRONT_R_MAX=params.front_r_max, FRONT_R_MIN=params.front_r_min,
        FRONT_C_MAX=params.front_c_max, FRONT_C_MIN=params.front_c_min, FRONT_WIDTH=params.front_width,
        FRONT_HEIGHT=params.front_height)
    ns = {"np": _NumpyCompat() if others else np, "numpy": np, "math": math, "cfg": cfg,
          "TOP_X_MIN": params.x_min, "TOP_X_MAX": params.x_max, "TOP_Y_MIN": params.y_min, "TOP_Y_MAX": params.y_max,
          "TOP_Z_MIN": params.z_min, "TOP_Z_MAX": params.z_max, "TOP_X_DIVISION": params.x_div,
          "TOP_Y_DIVISION": params.y_div, "TOP_Z_DIVISION": params.z_div}
    exec(compile(module, REFERENCE_DATA_PY, "exec"), ns)
    if others:
        front_slow = ns["lidar_to_front"]
        return ns["lidar_to_top_fast"], (lambda lidar: front_slow(None, lidar))
    front_fast = ns["lidar_to_front_fast"]
    return ns["lidar_to_top"], (lambda lidar: front_fast(None, lidar))
