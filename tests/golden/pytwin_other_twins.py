"""Executes the reference's two REMAINING Python twins -- `lidar_to_top_fast` (src/data.py:369-444) and
`Preprocess.lidar_to_front` (:56-111) -- out of /root/reference through the same AST loader as the two that are
pinned (tests/golden/pytwin_reference.py), and records how they relate to them:

    python tests/golden/pytwin_other_twins.py        # build container only; writes pytwin_other_twins.json

Neither is on any call path of the reference (`preprocess_for_mv3d.py` and the loaders call `lidar_to_top` and
`lidar_to_front_fast`); they are recorded so that nobody has to wonder whether a third semantics hides in them.
"""
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.dirname(HERE), os.path.dirname(os.path.dirname(HERE)), HERE]


def compare():
    import numpy as np
    import pytwin_reference as ref
    from mv3d_b200.params import PYTWIN_KITTI as p
    from mv3d_b200.sweeps import hdl64_sweep
    top_ref, front_ref = ref.load(p)
    top_fast, front_slow = ref.load(p, others=True)
    out = {"params": "PYTWIN_KITTI", "cases": {}}
    for name, pts in (("hdl64_0_every_8th", hdl64_sweep(0)[::8].copy()), ("hdl64_5_first_20000", hdl64_sweep(5)[:20000].copy())):
        a, b = top_ref(pts.copy()), top_fast(pts.copy())
        zn = p.zn
        b_flipped = b[::-1, ::-1]  # lidar_to_top writes top[-x-1, -y-1] (:345,:352), lidar_to_top_fast top[x, y] (:424-436)
        rec = {"points": int(pts.shape[0]), "top_shape_equal": list(a.shape) == list(b.shape),
               "top_equal_unflipped": bool(np.array_equal(a, b)),
               "heights_differ_after_flip": int(np.count_nonzero(a[..., :zn] != b_flipped[..., :zn])),
               "intensity_differs_after_flip": int(np.count_nonzero(a[..., zn] != b_flipped[..., zn])),
               "density_differs_after_flip": int(np.count_nonzero(a[..., zn + 1] != b_flipped[..., zn + 1])),
               "occupied_cells": int(np.count_nonzero(a[..., zn + 1]))}
        f, g = front_ref(pts.copy()), front_slow(pts.copy())
        d = np.abs(f.astype(np.float64) - g.astype(np.float64))
        rec.update({"front_shape_equal": list(f.shape) == list(g.shape),
                    "front_pixels_differ": int(np.count_nonzero((f != g).any(axis=2))),
                    "front_pixels_occupancy_differs": int(np.count_nonzero(f.any(axis=2) != g.any(axis=2))),
                    "front_max_abs_diff": float(d.max()), "front_occupied_pixels": int(np.count_nonzero(f.any(axis=2)))})
        out["cases"][name] = rec
    return out


if __name__ == "__main__":
    import pytwin_reference as ref
    if os.environ.get("NPY_DISABLE_CPU_FEATURES") is None:  # atan2f from libm, like the numpy of the reference's time
        raise SystemExit(subprocess.call([sys.executable, os.path.abspath(__file__)], env={**os.environ, **ref.LIBM_ENV}))
    res = compare()
    json.dump(res, open(os.path.join(HERE, "pytwin_other_twins.json"), "w"), indent=1)
    print(json.dumps(res, indent=1))
