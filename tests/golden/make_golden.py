"""Generates tests/golden/ver3_golden.npz by running the REFERENCE's own C code (oracle/_ref, compiled
unmodified from /root/reference/src/lidar_data_preprocess/Python_to_C_Interface/ver3/*.c by oracle/Makefile).

Run in the build container only (the GPU box has no /root/reference):  python tests/golden/make_golden.py
Outputs are stored sparsely (flat indices + values of non-zero cells) to keep the fixture small.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import cases  # noqa: E402


def golden_cases():
    out = cases.quirk_cases() + cases.small_cases()
    out += [c for c in cases.sweep_cases(frames=(0,)) if c[0] in ("hdl64_0", "hdl64_0_zround", "hdl64_0_didi")]
    return out


# Per-point pixel fixtures: the front view is a "last point wins" image, so whole-image comparisons of dense clouds
# hide single-point differences in the pixel index. These clouds are run through the reference build ONE POINT PER CALL
# (oracle/ref_probe.c) on a coarse bird's-eye-view grid that gates nothing out, and the reference's (cell, pixel) per
# point is what gets committed.
def probe_top():
    from mv3d_b200.params import TopParams
    return TopParams(-2000.0, 1000.0, -2000.0, 1000.0, -2000.0, 1000.0, 1000.0, 1000.0, 1000.0, 3, 3, 3)


def point_clouds():
    """name -> (N, 4) float32 cloud; seeded, regenerated identically by the tests."""
    from mv3d_b200.sweeps import hdl64_sweep
    rng = np.random.default_rng(20261105)
    n = 120000
    box = np.empty((n, 4), np.float32)  # the config-1 box with margins: every azimuth bin of +-45 deg and beyond
    box[:, 0] = rng.uniform(-5.0, 75.0, n)
    box[:, 1] = rng.uniform(-45.0, 45.0, n)
    box[:, 2] = rng.uniform(-4.5, 1.0, n)
    box[:, 3] = rng.random(n)
    ring = np.empty((60000, 4), np.float32)  # all four quadrants, near and far, steep and flat
    ang = rng.uniform(-np.pi, np.pi, 60000)
    rad = np.exp(rng.uniform(np.log(0.05), np.log(120.0), 60000))
    ring[:, 0] = rad * np.cos(ang)
    ring[:, 1] = rad * np.sin(ang)
    ring[:, 2] = rad * np.tan(rng.uniform(-0.5, 0.5, 60000))
    ring[:, 3] = rng.random(60000)
    axes = np.array([[1, 0, 0, .5], [-1, 0, 0, .5], [0, 1, 0, .5], [0, -1, 0, .5], [0, 0, 1, .5], [0, 0, -1, .5],
                     [1, 1, 0, .5], [1, -1, 0, .5], [-1, 1, 0, .5], [-1, -1, 0, .5], [0, 0, 0, .5],
                     [33.7921562, 5.01259804, -1.0, 0.5],  # round-1 advisor case: column 304, not 303
                     [1e-3, 1e-30, 1e-3, .5], [50, 1e-20, -1, .5], [50, -1e-20, -1, .5]], np.float32)
    return {"box": box, "ring": ring, "axes": axes, "hdl64_7": hdl64_sweep(7)[::2].copy()}


def pack(store, name, key, arr):
    flat = arr.reshape(-1)
    idx = np.flatnonzero((flat != 0) | np.isnan(flat))
    store[f"{name}/{key}/idx"] = idx.astype(np.int32)
    store[f"{name}/{key}/val"] = flat[idx]


def unpack(store, name, key, shape):
    out = np.zeros(int(np.prod(shape)), dtype=np.float32)
    out[store[f"{name}/{key}/idx"]] = store[f"{name}/{key}/val"]
    return out.reshape(shape)


def main():
    import oracle
    oracle.build()
    assert oracle.reference_available(), "needs /root/reference to build oracle/_ref"
    store = {}
    for name, pts, top, front in golden_cases():
        pack(store, name, "top", oracle.top_maps(pts, top, "reference"))
        t, f = oracle.top_and_front_maps(pts, top, front, "reference")
        pack(store, name, "topfront_top", t)
        pack(store, name, "topfront_front", f)
    path = os.path.join(HERE, "ver3_golden.npz")
    np.savez_compressed(path, **store)
    print("wrote", path, os.path.getsize(path), "bytes,", len(golden_cases()), "cases")
    from mv3d_b200.params import KITTI_FRONT_64x512
    keys = {}
    for name, pts in point_clouds().items():
        cell, pix = oracle.reference_point_keys(pts, probe_top(), KITTI_FRONT_64x512)
        keys[name + "/cell"] = cell.astype(np.int8)
        keys[name + "/pix"] = pix
    path = os.path.join(HERE, "ver3_point_keys.npz")
    np.savez_compressed(path, **keys)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
