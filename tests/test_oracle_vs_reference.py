"""CPU tests: pin the restated oracle (oracle/view_oracle.c) to the reference's own C code (oracle/_ref,
compiled unmodified from /root/reference) and to the golden fixtures that build produced."""
import os

import numpy as np
import pytest

import cases
from mv3d_b200.params import KITTI_FRONT_64x512, KITTI_PAPER_TOP

ALL_CASES = cases.quirk_cases() + cases.small_cases() + cases.sweep_cases() + cases.fuzz_cases()


def _needs_ref(oracle_mod):
    if not oracle_mod.reference_available():
        pytest.skip("oracle/_ref not built (reference sources absent and no prebuilt copy)")


@pytest.mark.parametrize("name,pts,top,front", ALL_CASES, ids=[c[0] for c in ALL_CASES])
def test_restated_equals_reference(oracle_mod, name, pts, top, front):
    _needs_ref(oracle_mod)
    a = oracle_mod.top_maps(pts, top, "reference")
    b = oracle_mod.top_maps(pts, top, "restated")
    assert np.array_equal(a, b, equal_nan=True), f"{name}: createTopMaps differs in {np.count_nonzero(a != b)} cells"
    at, af = oracle_mod.top_and_front_maps(pts, top, front, "reference")
    bt, bf = oracle_mod.top_and_front_maps(pts, top, front, "restated")
    assert np.array_equal(at, bt, equal_nan=True), f"{name}: top differs in {np.count_nonzero(at != bt)} cells"
    assert np.array_equal(af, bf, equal_nan=True), f"{name}: front differs in {np.count_nonzero(af != bf)} cells"


def test_known_answers(oracle_mod):
    """Hand-checked values for the quirks, on the restatement."""
    qc = {c[0]: c for c in cases.quirk_cases()}
    k = KITTI_PAPER_TOP
    # Q1
    top = oracle_mod.top_maps(qc["q1_float_rounding"][1], k)
    assert top[9, 9, 5] == np.float32(0.99999905) and top[9, 9, 6] == 0
    # Q2: truncation counts z in (z_MIN - z_DIV, z_MIN); floor does not
    t_trunc = oracle_mod.top_maps(qc["q2_below_zmin"][1], k)
    t_floor, _ = oracle_mod.top_and_front_maps(qc["q2_below_zmin"][1], k, KITTI_FRONT_64x512)
    lut = oracle_mod.density_table()
    assert t_trunc[19, 399, 26] == lut[3] and t_floor[19, 399, 26] == lut[1]
    assert t_trunc[19, 399, 0] == t_floor[19, 399, 0] != 0
    # Q3: last row / column never written
    t3 = oracle_mod.top_maps(qc["q3_edges"][1], k)
    assert not t3[-1].any() and not t3[:, -1].any()
    assert t3[702, 798, 26] != 0 and t3[0, 0, 26] != 0
    # Q4
    t4 = oracle_mod.top_maps(qc["q4_z_bounds"][1], k)
    assert t4[19, 419, 26] == lut[1] and not t4[19, 419, :26].any()
    # Q5: first of the tied maxima wins
    t5 = oracle_mod.top_maps(qc["q5_ties"][1], k)
    assert t5[49, 409, 25] == np.float32(0.5) and t5[59, 409, 25] == np.float32(0.4)
    # Q6: density table and saturation
    t6 = oracle_mod.top_maps(qc["q6_density"][1], k)
    for c in (1, 2, 30, 31, 40):
        assert t6[int(np.float32((np.float32(10.05 + 0.1 * c)) / np.float32(0.1))) - 1, 429, 26] == (lut[c] if c < 32 else 1.0)
    assert lut[31] == 1.0 and lut[0] == 0.0 and abs(lut[1] - 0.2) < 1e-7
    # Q8: last in-range point wins, truncated integers
    _, f8 = oracle_mod.top_and_front_maps(qc["q8_front_last"][1], k, KITTI_FRONT_64x512)
    nz = np.argwhere(f8.any(axis=2))
    assert len(nz) == 1
    r, c = nz[0]
    assert tuple(f8[r, c]) == (163.0, 42.0, 102.0)


def test_density_model_is_bit_exact(oracle_mod):
    """Q6: (float)((double)logf(c+1) / log(32)) -- 1 ulp away from an all-double model for some counts."""
    lut = oracle_mod.density_table()
    c = np.arange(32, dtype=np.float32)
    model = (np.log(c + np.float32(1)).astype(np.float32).astype(np.float64) / np.log(32.0)).astype(np.float32)
    assert np.array_equal(np.minimum(model, np.float32(1)), lut)


GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "ver3_golden.npz")


def test_restated_equals_golden(oracle_mod):
    """Outputs of the reference build, committed as fixtures (tests/golden/make_golden.py), so the pin
    also holds on machines without /root/reference."""
    g = np.load(GOLDEN)
    import golden.make_golden as mg
    for name, pts, top, front in mg.golden_cases():
        t = oracle_mod.top_maps(pts, top)
        tf, ff = oracle_mod.top_and_front_maps(pts, top, front)
        for key, arr in (("top", t), ("topfront_top", tf), ("topfront_front", ff)):
            assert np.array_equal(mg.unpack(g, name, key, arr.shape), arr, equal_nan=True), f"{name}/{key}"


POINT_KEYS = os.path.join(os.path.dirname(__file__), "golden", "ver3_point_keys.npz")


def test_restated_point_keys_equal_golden(oracle_mod):
    """Per-point front-view pixel and bird's-eye-view cell of the REFERENCE build, one point per call (committed by
    tests/golden/make_golden.py): whole-image comparisons of dense clouds hide single-point differences, this does
    not. Pins the types of LidarTopAndFrontPreprocess.c:173-174 (row bin in double, column bin in float/atan2f)."""
    import golden.make_golden as mg
    from mv3d_b200.params import KITTI_FRONT_64x512
    g = np.load(POINT_KEYS)
    for name, pts in mg.point_clouds().items():
        cell, pix = oracle_mod.point_keys(pts, mg.probe_top(), KITTI_FRONT_64x512)
        assert np.array_equal(cell, g[name + "/cell"].astype(np.int32)), name
        bad = np.flatnonzero(pix != g[name + "/pix"])
        assert bad.size == 0, f"{name}: {bad.size} points land in another pixel, first {pts[bad[0]]}"
    # the round-1 advisor's counter-example: (28, 304), not (28, 303)
    p = np.array([[33.7921562, 5.01259804, -1.0, 0.5]], np.float32)
    assert oracle_mod.point_keys(p, mg.probe_top(), KITTI_FRONT_64x512)[1][0] == 28 * 512 + 304


def test_restated_point_keys_equal_reference(oracle_mod):
    """The same comparison against the reference build itself, on fresh random points (needs oracle/_ref)."""
    _needs_ref(oracle_mod)
    import golden.make_golden as mg
    from mv3d_b200.params import KITTI_FRONT_64x512, FrontParams
    rng = np.random.default_rng()
    n = 150000
    pts = np.empty((n, 4), np.float32)
    pts[:, 0] = rng.uniform(-10.0, 80.0, n)
    pts[:, 1] = rng.uniform(-50.0, 50.0, n)
    pts[:, 2] = rng.uniform(-5.0, 2.0, n)
    pts[:, 3] = 0.5
    for front in (KITTI_FRONT_64x512, FrontParams(-20.0, 20.0, -45.0, 45.0, 0.1, 0.2)):
        rc, rp = oracle_mod.reference_point_keys(pts[: n if front is KITTI_FRONT_64x512 else 20000], mg.probe_top(), front)
        oc, op = oracle_mod.point_keys(pts[: len(rc)], mg.probe_top(), front)
        assert np.array_equal(rc, oc)
        bad = np.flatnonzero(rp != op)
        assert bad.size == 0, f"{bad.size} points differ, first {pts[bad[0]]} ref {rp[bad[0]]} restated {op[bad[0]]}"
