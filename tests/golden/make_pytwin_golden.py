"""Generates tests/golden/pytwin_golden.npz by EXECUTING the reference's own Python encoders (pytwin_reference.py
compiles `lidar_to_top` and `Preprocess.lidar_to_front_fast` out of /root/reference/src/data.py at run time).

Run in the build container only:  python tests/golden/make_pytwin_golden.py
It re-executes itself with NPY_DISABLE_CPU_FEATURES set so that float32 np.arctan2 comes from the C library's atan2f,
as it did in the numpy the reference was written against. Outputs are stored sparsely.
"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import pytwin_reference as ref  # noqa: E402

if os.environ.get("NPY_DISABLE_CPU_FEATURES") != ref.LIBM_ENV["NPY_DISABLE_CPU_FEATURES"]:
    os.execve(sys.executable, [sys.executable] + sys.argv, {**os.environ, **ref.LIBM_ENV})

import numpy as np  # noqa: E402
import pytwin_cases  # noqa: E402
from make_golden import pack  # noqa: E402


def main():
    assert ref.available(), "needs /root/reference"
    store = {}
    for name, pts, p in pytwin_cases.all_cases():
        top, front = ref.load(p)
        pack(store, name, "top", top(pts.copy()))
        try:
            f = front(pts.copy())
        except ValueError:  # the reference raises when no point survives the crop; the new path returns zeros
            assert name == "empty"
            f = np.zeros(p.front_shape, np.float32)
        pack(store, name, "front", f)
        print(name, "done", flush=True)
    path = os.path.join(HERE, "pytwin_golden.npz")
    np.savez_compressed(path, **store)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
