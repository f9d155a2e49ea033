"""View-encoder parameter sets.

The reference passes every range/resolution value explicitly through its C ABI
(`ver3/LidarTopPreprocess.c:30`, `ver3/LidarTopAndFrontPreprocess.c:30`); the caller derives the
grid sizes (`src/data.py:270-272` with `math.floor`, `ver3/ExampleUsage_TopAndFront.py:134-140`
with `int()`). These dataclasses hold one such parameter set and reproduce that derivation.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, replace


@dataclass(frozen=True)
class TopParams:
    """Bird's-eye-view box and resolution: the 9 floats + 3 ints of `createTopMaps`."""

    x_min: float
    x_max: float
    y_min: float
    y_max: float
    z_min: float
    z_max: float
    x_div: float
    y_div: float
    z_div: float
    xn: int = -1
    yn: int = -1
    zn: int = -1
    # What `clidar_to_top` (src/data.py:265-291) does around the C call, fused into the encode (MV3D_TOP_* flags):
    flip: bool = False               # np.flipud(np.fliplr(top_flip)), :290
    filter_center_car: bool = False  # filter_center_car, :224-227, for the didi data sets (:266-267)

    @property
    def flags(self) -> int:
        return (1 if self.flip else 0) | (2 if self.filter_center_car else 0)

    def __post_init__(self):
        # Grid sizes as the reference caller computes them, in double (src/data.py:270-272).
        if self.xn < 0:
            object.__setattr__(self, "xn", math.floor((self.x_max - self.x_min) / self.x_div))
        if self.yn < 0:
            object.__setattr__(self, "yn", math.floor((self.y_max - self.y_min) / self.y_div))
        if self.zn < 0:
            object.__setattr__(self, "zn", math.floor((self.z_max - self.z_min) / self.z_div))

    @property
    def shape(self):
        """Output layout `[Xn][Yn][Zn+2]`, channel-minor (LidarTopPreprocess.c:200)."""
        return (self.xn, self.yn, self.zn + 2)

    @property
    def floats(self):
        return (self.x_min, self.x_max, self.y_min, self.y_max, self.z_min, self.z_max,
                self.x_div, self.y_div, self.z_div)

    def as_top_paras(self):
        """The 12-tuple `ver3/createTopMaps.py:5-8` unpacks."""
        return (*self.floats, self.xn, self.yn, self.zn)


@dataclass(frozen=True)
class FrontParams:
    """Front-view (cylindrical) window: the 6 floats + 3 sizes of `createTopAndFrontMaps`.

    `phi_min/phi_max/theta_min/theta_max` are accepted and ignored by the reference routine
    (LidarTopAndFrontPreprocess.c:30 vs :173-179); they only size the image on the caller side.
    """

    phi_min: float
    phi_max: float
    theta_min: float
    theta_max: float
    phi_div: float
    theta_div: float
    rn: int = -1
    cn: int = -1
    fn: int = 3

    def __post_init__(self):
        if self.rn < 0:  # ver3/ExampleUsage_TopAndFront.py:138-139
            object.__setattr__(self, "rn", int((self.phi_max - self.phi_min) / self.phi_div))
        if self.cn < 0:
            object.__setattr__(self, "cn", int((self.theta_max - self.theta_min) / self.theta_div))

    @property
    def shape(self):
        """Output layout `[R][C][F]` (LidarTopAndFrontPreprocess.c:185)."""
        return (self.rn, self.cn, self.fn)

    def as_front_paras(self):
        """The 9-tuple `ver3/createTopAndFrontMaps.py:10-13` unpacks (note: Rn before Cn)."""
        return (self.phi_min, self.phi_max, self.theta_min, self.theta_max,
                self.phi_div, self.theta_div, self.rn, self.cn, self.fn)


# BASELINE.json configs 1/2/4: MV3D-paper KITTI crop, 704 x 800 x (25+2) at 0.1 m.
KITTI_PAPER_TOP = TopParams(0.0, 70.4, -40.0, 40.0, -4.2, 0.8, 0.1, 0.1, 0.2, 704, 800, 25)
# Front view 64 beams x 512 azimuth bins: phi +-13.4 deg, theta +-45 deg.
KITTI_FRONT_64x512 = FrontParams(-13.4, 13.4, -45.0, 45.0, 26.8 / 64, 90.0 / 512, 64, 512, 3)
# BASELINE.json config 5: dense aggregated cloud, 1408 x 1600 x (8+2) at 0.05 m.
DENSE_TOP = TopParams(0.0, 70.4, -40.0, 40.0, -2.0, 1.2, 0.05, 0.05, 0.4, 1408, 1600, 8)
# The reference's own presets (src/config.py:154-187).
REFERENCE_KITTI_TOP = TopParams(0.0, 80.0, -30.0, 30.0, -4.2, 0.8, 0.1, 0.1, 0.2)
REFERENCE_DIDI_TOP = TopParams(-45.0, 45.0, -10.0, 10.0, -3.0, 0.7, 0.2, 0.2, 0.3)
REFERENCE_DIDI2_TOP = TopParams(-50.0, 50.0, -30.0, 30.0, -3.5, 0.6, 0.2, 0.2, 0.3)

# `clidar_to_top` per data set (src/data.py:265-291): the C routine, flipped; the centre-car filter only for `didi`
# (and `test`) -- unlike the Python encoder, which also filters `didi2` (:313).
CLIDAR_KITTI = replace(REFERENCE_KITTI_TOP, flip=True)
CLIDAR_DIDI = replace(REFERENCE_DIDI_TOP, flip=True, filter_center_car=True)
CLIDAR_DIDI2 = replace(REFERENCE_DIDI2_TOP, flip=True)
CLIDAR_PRESETS = {"kitti": CLIDAR_KITTI, "didi": CLIDAR_DIDI, "test": CLIDAR_DIDI, "didi2": CLIDAR_DIDI2}

PRESETS = {
    "kitti_paper": KITTI_PAPER_TOP,
    "dense": DENSE_TOP,
    "kitti": REFERENCE_KITTI_TOP,
    "didi": REFERENCE_DIDI_TOP,
    "didi2": REFERENCE_DIDI2_TOP,
}

__all__ = ["TopParams", "FrontParams", "KITTI_PAPER_TOP", "KITTI_FRONT_64x512", "DENSE_TOP",
           "REFERENCE_KITTI_TOP", "REFERENCE_DIDI_TOP", "REFERENCE_DIDI2_TOP", "PRESETS", "replace",
           "CLIDAR_KITTI", "CLIDAR_DIDI", "CLIDAR_DIDI2", "CLIDAR_PRESETS"]


@dataclass(frozen=True)
class PyTwinParams:
    """Parameter set of the reference's PYTHON encoders (`src/data.py:113-180` `lidar_to_front_fast`, `:296-367`
    `lidar_to_top`): the module-level `TOP_*` constants of `src/config.py:154-189` plus the front-view constants of
    `src/config.py:32-42`. Values are Python floats (doubles), as in the reference; the encoders narrow them to
    float32 where numpy does."""

    x_min: float
    x_max: float
    y_min: float
    y_max: float
    z_min: float
    z_max: float
    x_div: float
    y_div: float
    z_div: float
    filter_center_car: bool = False            # src/data.py:313-314 (didi / didi2 / test data sets only)
    ang_res: float = 0.08 / 180 * math.pi      # VELODYNE_ANGULAR_RESOLUTION
    vert_res: float = 0.4 / 180 * math.pi      # VELODYNE_VERTICAL_RESOLUTION
    velodyne_height: float = 1.73
    front_r_offset: int = 70
    front_c_offset: int = 750
    front_r_min: int = -70
    front_r_max: int = 30
    front_c_min: int = -750
    front_c_max: int = 750
    front_width: int = 1500
    front_height: int = 100

    @property
    def xn(self):
        """src/data.py:327: `int((TOP_X_MAX-TOP_X_MIN)//TOP_X_DIVISION)+1`, in double."""
        return int((self.x_max - self.x_min) // self.x_div) + 1

    @property
    def yn(self):
        return int((self.y_max - self.y_min) // self.y_div) + 1

    @property
    def zn(self):
        """src/data.py:329: true division, truncated."""
        return int((self.z_max - self.z_min) / self.z_div)

    @property
    def top_shape(self):
        return (self.xn, self.yn, self.zn + 2)

    @property
    def front_shape(self):
        """`[FRONT_WIDTH][FRONT_HEIGHT][3]`, indexed `[c][r]` (src/data.py:167)."""
        return (self.front_width, self.front_height, 3)


# The reference's data-set presets (src/config.py:154-187) for the Python encoders.
PYTWIN_KITTI = PyTwinParams(0, 80, -30, 30, -4.2, 0.8, 0.1, 0.1, 0.2)
PYTWIN_DIDI = PyTwinParams(-45, 45, -10, 10, -3.0, 0.7, 0.2, 0.2, 0.3, filter_center_car=True)
PYTWIN_DIDI2 = PyTwinParams(-50, 50, -30, 30, -3.5, 0.6, 0.2, 0.2, 0.3, filter_center_car=True)
PYTWIN_PRESETS = {"kitti": PYTWIN_KITTI, "didi": PYTWIN_DIDI, "didi2": PYTWIN_DIDI2}
