"""Synthetic KITTI-shaped Velodyne sweeps (there is no network and the reference ships no point data).

The on-disk record the reference reads is `np.fromfile(path, float32).reshape(-1, 4)` =
`x, y, z, intensity` (`src/raw_data.py:137-139`, `src/utils/batch_loading.py:656`); these generators
produce arrays of that exact layout. Recipe: SURVEY.md section 8(d).
"""
from __future__ import annotations

import numpy as np

HDL64_BEAMS = 64
HDL64_AZIMUTH_STEPS = 1875  # 64 * 1875 = 120 000 returns per sweep
SENSOR_HEIGHT_M = 1.73
MAX_RANGE_M = 80.0


def hdl64_sweep(frame_id: int = 0, beams: int = HDL64_BEAMS, azimuth_steps: int = HDL64_AZIMUTH_STEPS) -> np.ndarray:
    """One HDL-64E-like sweep, `(beams*azimuth_steps, 4)` float32, beam-major point order.

    Beam elevations run linearly from +2.0 deg to -24.8 deg; downward beams hit a flat ground plane
    1.73 m below the sensor (capped at 80 m), the rest return at 80 m; every return is pulled in by a
    U(0.6, 1.0) factor (obstacle clutter); intensity is U[0, 1). Seeded by `1000 + frame_id`.
    """
    rng = np.random.default_rng(1000 + int(frame_id))
    elev = np.deg2rad(np.linspace(2.0, -24.8, beams))
    azim = np.linspace(0.0, 2.0 * np.pi, azimuth_steps, endpoint=False)
    with np.errstate(divide="ignore"):
        ground = np.where(elev < 0.0, SENSOR_HEIGHT_M / np.tan(-elev), MAX_RANGE_M)
    rng_range = np.minimum(ground, MAX_RANGE_M)[:, None] * rng.uniform(0.6, 1.0, size=(beams, azimuth_steps))
    ce, se = np.cos(elev)[:, None], np.sin(elev)[:, None]
    pts = np.empty((beams, azimuth_steps, 4), dtype=np.float32)
    pts[..., 0] = rng_range * ce * np.cos(azim)[None, :]
    pts[..., 1] = rng_range * ce * np.sin(azim)[None, :]
    pts[..., 2] = rng_range * se
    pts[..., 3] = rng.random(size=(beams, azimuth_steps), dtype=np.float32)
    return np.ascontiguousarray(pts.reshape(-1, 4))


def dense_cloud(num_points: int = 2_000_000, copies: int = 17, shift_m: float = 3.0 / 16.0, frame_id: int = 0) -> np.ndarray:
    """BASELINE.json config 5: `copies` sweeps, copy j shifted by j*shift_m in x, truncated to num_points."""
    parts = []
    for j in range(copies):
        s = hdl64_sweep(frame_id + j).copy()
        s[:, 0] += np.float32(j * shift_m)
        parts.append(s)
    return np.ascontiguousarray(np.concatenate(parts, axis=0)[:num_points])


def uniform_in_box(num_points: int, top, seed: int = 0, margin: float = 0.0) -> np.ndarray:
    """Uniform random points over (a slightly inflated copy of) the BEV box: worst case for the gate
    (every point kept) and, with `margin` > 0, exercises every rejection edge."""
    rng = np.random.default_rng(seed)
    lo = np.array([top.x_min - margin, top.y_min - margin, top.z_min - margin], dtype=np.float64)
    hi = np.array([top.x_max + margin, top.y_max + margin, top.z_max + margin], dtype=np.float64)
    pts = np.empty((num_points, 4), dtype=np.float32)
    pts[:, :3] = rng.uniform(lo, hi, size=(num_points, 3))
    pts[:, 3] = rng.random(num_points, dtype=np.float32)
    return pts


def pack_sweeps(sweeps):
    """Concatenate ragged sweeps into one `(total, 4)` float32 array + `(batch+1,)` int32 offsets
    (the layout the batched device entry point consumes)."""
    offsets = np.zeros(len(sweeps) + 1, dtype=np.int32)
    for i, s in enumerate(sweeps):
        offsets[i + 1] = offsets[i] + s.shape[0]
    if len(sweeps) == 0:
        return np.zeros((0, 4), np.float32), offsets
    return np.ascontiguousarray(np.concatenate(sweeps, axis=0), dtype=np.float32), offsets
