"""mv3d_b200 -- B200-native LiDAR view encoder behind MV3D's own ctypes/.so boundary.

Only the hot path lives here: `csrc/` (sm_100a kernels + the C ABI of include/mv3d_encode.h) and the
host-side mirror of the reference's wrappers. Nothing in this package imports `oracle/`.
"""
from .params import (DENSE_TOP, KITTI_FRONT_64x512, KITTI_PAPER_TOP, PRESETS, REFERENCE_KITTI_TOP,
                     FrontParams, TopParams)

__all__ = ["TopParams", "FrontParams", "KITTI_PAPER_TOP", "KITTI_FRONT_64x512", "DENSE_TOP",
           "REFERENCE_KITTI_TOP", "PRESETS"]
__version__ = "0.1.0"
