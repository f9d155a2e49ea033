"""Shared cases of the Python-twin parity tests: (name, points, PyTwinParams). Hand-built clouds exercise every
difference from the C routine listed in SURVEY.md section 8(f) row 1."""
import numpy as np

from mv3d_b200.params import PYTWIN_DIDI, PYTWIN_DIDI2, PYTWIN_KITTI, PyTwinParams
from mv3d_b200.sweeps import hdl64_sweep

SMALL = PyTwinParams(0, 8, -3, 3, -1.0, 1.0, 0.2, 0.2, 0.4)
SMALL_FILTER = PyTwinParams(-8, 8, -3, 3, -1.0, 1.0, 0.2, 0.2, 0.4, filter_center_car=True)


def P(*rows):
    return np.array(rows, dtype=np.float32).reshape(-1, 4)


def quirk_cases():
    k = PYTWIN_KITTI
    cases = []
    # open-interval crop: points exactly on a bound are dropped, one float32 step inside is kept
    e = np.float32
    cases.append(("crop_edges", P([0.0, 0.0, -1.0, 0.1], [np.nextafter(e(0), e(1)), 0.0, -1.0, 0.2], [80.0, 1.0, -1.0, 0.3],
                                  [np.nextafter(e(80), e(0)), 1.0, -1.0, 0.4], [5.0, -30.0, -1.0, 0.5],
                                  [5.0, np.nextafter(e(30), e(0)), -1.0, 0.6], [5.0, 5.0, -4.2, 0.7],
                                  [5.0, 5.0, np.nextafter(e(-4.2), e(0)), 0.8], [6.0, 5.0, 0.8, 0.9],
                                  [6.0, 5.0, np.nextafter(e(0.8), e(0)), 0.95]), k))
    # floor_divide is numpy's fmod-based one, not floor(a/b): multiples of the cell size land differently
    xs = np.arange(1, 400, dtype=np.float32) * np.float32(0.1)
    cases.append(("floor_divide", np.stack([xs, xs * 0 + 1.0, xs * 0 - 1.0, xs * 0 + 0.5], axis=1).astype(np.float32), k))
    # a point whose qz is integral closes the slice below (value 1.0) and opens its own (value 0)
    zs = np.float32(-4.2) + np.arange(0, 26, dtype=np.float32) * np.float32(0.2)
    rows = [[10.05 + 0.1 * i, 2.05, z, 0.3 + 0.01 * i] for i, z in enumerate(zs)]
    rows += [[10.05 + 0.1 * i, 4.05, np.nextafter(np.float32(z), np.float32(9)), 0.5] for i, z in enumerate(zs)]
    cases.append(("integral_qz", P(*rows), k))
    # intensity: first point in input order among equal largest qz; density table incl. saturation
    rows = []
    for c in range(1, 41):
        for j in range(c):
            rows.append([20.05 + 0.1 * c, -3.05, -1.0 if j % 3 == 0 else -2.0 - 0.01 * j, 0.01 * (j + 1)])
    cases.append(("ties_density", P(*rows), k))
    # front view: several points in one pixel (float32 running mean in input order), window and offset edges
    rng = np.random.default_rng(3)
    blob = np.empty((600, 4), np.float32)
    blob[:, 0] = 20.0 + rng.uniform(0, 0.02, 600)
    blob[:, 1] = 1.0 + rng.uniform(0, 0.02, 600)
    blob[:, 2] = -1.0 + rng.uniform(0, 0.02, 600)
    blob[:, 3] = rng.random(600)
    steep = P([2.0, 0.1, 0.79, 0.5], [1.0, 0.0001, 0.75, 0.6], [60.0, 29.0, -4.0, 0.7], [60.0, -29.0, 0.5, 0.8],
              [0.001, 0.0005, -0.5, 0.9], [3.0, 2.9, -3.9, 0.2])
    cases.append(("front_pixels", np.concatenate([blob, steep]), k))
    # non-finite input never passes the crop
    nan, inf = np.nan, np.inf
    cases.append(("nonfinite", P([nan, 1, -1, 0.5], [1, nan, -1, 0.5], [1, 1, nan, 0.5], [inf, 1, -1, 0.5], [1, -inf, -1, 0.5],
                                 [5.0, 5.0, -1.0, nan], [6.0, 5.0, -1.0, 0.5]), k))
    return cases


def small_cases():
    rng = np.random.default_rng(11)
    pts = np.empty((5000, 4), np.float32)
    pts[:, 0] = rng.uniform(-9, 9, 5000)
    pts[:, 1] = rng.uniform(-3.5, 3.5, 5000)
    pts[:, 2] = np.round(rng.uniform(-1.2, 1.2, 5000), 1)   # many integral qz and ties
    pts[:, 3] = rng.random(5000)
    return [("empty", np.zeros((0, 4), np.float32), SMALL), ("small_random", pts, SMALL),
            ("small_center_car", pts, SMALL_FILTER)]


def sweep_cases():
    s0 = hdl64_sweep(0)
    perm = np.random.default_rng(42).permutation(len(s0))
    return [("hdl64_0_kitti", s0, PYTWIN_KITTI), ("hdl64_0_permuted", np.ascontiguousarray(s0[perm]), PYTWIN_KITTI),
            ("hdl64_1_didi2", hdl64_sweep(1), PYTWIN_DIDI2), ("hdl64_2_didi", hdl64_sweep(2), PYTWIN_DIDI)]


def all_cases():
    return quirk_cases() + small_cases() + sweep_cases()
