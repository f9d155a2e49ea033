"""CPU tests: pin the Python-twin oracle (oracle/pytwin.py, oracle/pytwin_oracle.c) to the reference's own Python
encoders -- executed from /root/reference where it exists, and through the committed outputs of that run
(tests/golden/pytwin_golden.npz) everywhere."""
import os
import subprocess
import sys

import numpy as np
import pytest

import pytwin_cases
from golden.make_golden import unpack

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "pytwin_golden.npz")


@pytest.fixture(scope="module")
def pytwin():
    from oracle import pytwin
    pytwin.build()
    return pytwin


@pytest.mark.parametrize("name,pts,p", pytwin_cases.all_cases(), ids=[c[0] for c in pytwin_cases.all_cases()])
def test_restatement_equals_golden(pytwin, name, pts, p):
    g = np.load(GOLDEN)
    top = pytwin.lidar_to_top(pts, p)
    front = pytwin.lidar_to_front_fast(pts, p)
    assert top.shape == p.top_shape and front.shape == p.front_shape
    assert np.array_equal(unpack(g, name, "top", top.shape), top, equal_nan=True), f"{name} top"
    assert np.array_equal(unpack(g, name, "front", front.shape), front, equal_nan=True), f"{name} front"


def test_restatement_equals_executed_reference(pytwin):
    """Runs the reference's functions in a child process with numpy's SIMD dispatch off (atan2f from libm)."""
    from golden import pytwin_reference as ref
    if not ref.available():
        pytest.skip("/root/reference not present")
    code = r'''
import sys
sys.path[:0] = [%r, %r, %r]
import numpy as np
import pytwin_cases
from golden import pytwin_reference as ref
from oracle import pytwin
bad = 0
for name, pts, p in pytwin_cases.quirk_cases() + pytwin_cases.small_cases()[1:]:
    top, front = ref.load(p)
    bad += 0 if np.array_equal(top(pts.copy()), pytwin.lidar_to_top(pts, p), equal_nan=True) else 1
    bad += 0 if np.array_equal(front(pts.copy()), pytwin.lidar_to_front_fast(pts, p), equal_nan=True) else 1
print("MISMATCH", bad)
''' % (os.path.dirname(__file__), os.path.dirname(os.path.dirname(__file__)), os.path.join(os.path.dirname(__file__), "golden"))
    out = subprocess.run([sys.executable, "-c", code], env={**os.environ, **ref.LIBM_ENV}, capture_output=True, text=True)
    assert "MISMATCH 0" in out.stdout, out.stdout + out.stderr


def test_atan2f_port_equals_libm(pytwin):
    """The fdlibm restatement the CUDA path mirrors == this C library's atan2f, bit for bit."""
    rng = np.random.default_rng(0)
    for scale in (80.0, 1.0, 1e-3, 1e6):
        x = rng.uniform(-scale, scale, 2_000_000).astype(np.float32)
        y = rng.uniform(-scale, scale, 2_000_000).astype(np.float32)
        a, b = pytwin.atan2f(y, x, "fdlibm"), pytwin.atan2f(y, x, "libm")
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    x = (rng.standard_normal(2_000_000) * 10.0 ** rng.integers(-20, 20, 2_000_000)).astype(np.float32)
    y = (rng.standard_normal(2_000_000) * 10.0 ** rng.integers(-20, 20, 2_000_000)).astype(np.float32)
    x[:10] = [0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, 0.0, 5.0, -5.0, 1.0]
    y[:10] = [0.0, 0.0, 0.0, -0.0, 1.0, np.inf, -3.0, np.inf, -np.inf, 7.0]
    a, b = pytwin.atan2f(y, x, "fdlibm"), pytwin.atan2f(y, x, "libm")
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_floor_divide_restatement_equals_numpy(pytwin):
    rng = np.random.default_rng(1)
    for b in (0.1, 0.2, 0.05, 0.3, 0.4):
        a = rng.uniform(-5, 120, 1_000_000).astype(np.float32)
        a[::5] = np.round(a[::5], 1)
        got = pytwin.floor_divide_f32(a, b)
        assert np.array_equal(got, a // np.float32(b))
        # and it is NOT floor(a / b): the two differ on multiples of the cell size
    a = np.arange(1, 400, dtype=np.float32) * np.float32(0.1)
    assert np.count_nonzero(np.floor(a / np.float32(0.1)) != a // np.float32(0.1)) > 0


def test_known_answers(pytwin):
    from mv3d_b200.params import PYTWIN_KITTI as k
    qc = {c[0]: c for c in pytwin_cases.quirk_cases()}
    # crop: only the points one float32 step inside the box survive
    top = pytwin.lidar_to_top(qc["crop_edges"][1], k)
    assert np.count_nonzero(top[..., k.zn + 1]) == 5
    # output is written flipped: the point at x -> 0+, y = 0 lands in the last row
    assert top[k.xn - 1, k.yn - 1 - int(np.float32(30) // np.float32(0.1)), k.zn + 1] > 0
    # an integral qz closes the slice below with 1.0
    t = pytwin.lidar_to_top(qc["integral_qz"][1], k)
    cell = t[k.xn - 1 - 105, k.yn - 1 - 320]  # qz = 5 exactly: slice 4 holds 1.0, slice 5 holds 0
    assert cell[4] == 1.0 and cell[5] == 0.0 and not cell[:4].any()
    # density saturates at 31 points; intensity is the first of the tied maxima
    lut = pytwin.density_table()
    t = pytwin.lidar_to_top(qc["ties_density"][1], k)
    col = k.yn - 1 - int(np.float32(-3.05 + 30) // np.float32(0.1))
    for c in (1, 2, 30, 31, 40):
        row = k.xn - 1 - int(np.float32(20.05 + 0.1 * c) // np.float32(0.1))
        assert t[row, col, k.zn + 1] == lut[min(c, 31)]
        assert t[row, col, k.zn] == np.float32(0.01)
    assert lut[31] == 1.0 and abs(lut[1] - 0.2) < 1e-7


def test_other_two_twins_are_recorded():
    """`lidar_to_top_fast` (src/data.py:369-444) and `Preprocess.lidar_to_front` (:56-111) are on no call path of the
    reference; tests/golden/pytwin_other_twins.py executes them and records how they relate to the pinned pair:
    the first IS `lidar_to_top` without the x/y flip (every plane equal after flipping), the second is a different
    image from `lidar_to_front_fast` (double `math.atan2`, running mean) and is not reproduced. Where /root/reference
    exists the record is re-derived and must match."""
    import json
    here = os.path.join(os.path.dirname(__file__), "golden")
    rec = json.load(open(os.path.join(here, "pytwin_other_twins.json")))
    for case in rec["cases"].values():
        assert case["top_shape_equal"] and not case["top_equal_unflipped"]
        assert case["heights_differ_after_flip"] == case["intensity_differs_after_flip"] == case["density_differs_after_flip"] == 0
        assert case["occupied_cells"] > 1000
        assert case["front_shape_equal"] and case["front_pixels_differ"] > 0
    from golden import pytwin_reference as ref
    if not ref.available():
        return
    code = ("import sys, json; sys.path.insert(0, %r); import pytwin_other_twins as m; print('REC' + json.dumps(m.compare()))" % here)
    out = subprocess.run([sys.executable, "-c", code], env={**os.environ, **ref.LIBM_ENV}, capture_output=True, text=True)
    line = [l for l in out.stdout.splitlines() if l.startswith("REC")]
    assert line, out.stdout + out.stderr
    assert json.loads(line[0][3:]) == rec
