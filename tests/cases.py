"""Shared parity cases: seeded sweeps and hand-built known-answer clouds for every quirk of the
reference routine listed in SURVEY.md section 8(a) (Q1-Q12). Each case is (name, points, top, front)."""
import numpy as np

from mv3d_b200.params import (DENSE_TOP, KITTI_FRONT_64x512, KITTI_PAPER_TOP, REFERENCE_DIDI_TOP,
                              REFERENCE_KITTI_TOP, FrontParams, TopParams)
from mv3d_b200.sweeps import dense_cloud, hdl64_sweep, uniform_in_box

# A small grid so dense outputs stay tiny: 20 x 24 x (5+2), odd plane count like the real one.
SMALL_TOP = TopParams(0.0, 4.0, -2.4, 2.4, -1.0, 1.0, 0.2, 0.2, 0.4, 20, 24, 5)
SMALL_FRONT = FrontParams(-20.0, 20.0, -45.0, 45.0, 2.5, 5.0, 16, 18, 3)
# 1-cell-wide and degenerate grids
THIN_TOP = TopParams(0.0, 2.0, -1.0, 1.0, -1.0, 1.0, 1.0, 1.0, 1.0, 2, 2, 2)


def P(*rows):
    return np.array(rows, dtype=np.float32).reshape(-1, 4)


def quirk_cases():
    k = KITTI_PAPER_TOP
    cases = []
    # Q1: two separately rounded float ops: (-3.0 - -4.2)/0.2 = 5.99999905 -> slice 5, value 0.99999905
    cases.append(("q1_float_rounding", P([1.05, -39.0, -3.0, 0.25]), k, KITTI_FRONT_64x512))
    # Q2: z in (z_MIN - z_DIV, z_MIN): truncation keeps it for density, floor rejects it
    cases.append(("q2_below_zmin", P([2.05, 0.05, -4.25, 0.5], [2.05, 0.05, -4.0, 0.7], [2.05, 0.05, -4.39, 0.1],
                                     [2.05, 0.05, -4.41, 0.1]), k, KITTI_FRONT_64x512))
    # Q3: X == 0 / Y == 0 cells dropped, output shifted by -1, last row/column stay zero
    cases.append(("q3_edges", P([0.05, 0.0, -1.0, 0.3], [0.15, -39.95, -1.0, 0.4], [0.15, -39.85, -1.0, 0.5],
                                [70.35, 39.95, -1.0, 0.6], [70.39, 39.99, 0.0, 0.7], [70.4, 0.0, -1.0, 0.8],
                                [10.0, 40.0, -1.0, 0.9], [-0.05, 0.0, -1.0, 0.2], [10.0, -40.05, -1.0, 0.2]),
                  k, KITTI_FRONT_64x512))
    # Q4: z == z_MAX dropped; z == z_MIN kept and counted but h == 0 is not > 0
    cases.append(("q4_z_bounds", P([2.05, 2.05, 0.8, 0.5], [2.05, 2.05, -4.2, 0.6], [3.05, 2.05, 0.79999, 0.7]),
                  k, KITTI_FRONT_64x512))
    # Q5: equal heights -> the first point in input order supplies the intensity (also across slices)
    cases.append(("q5_ties", P([5.05, 1.05, -1.0, 0.5], [5.07, 1.07, -1.0, 0.9], [5.02, 1.02, -1.0, 0.1],
                               [6.05, 1.05, -2.0, 0.3], [6.05, 1.05, -1.5, 0.4], [6.05, 1.05, -1.5, 0.8],
                               [6.05, 1.05, -3.0, 0.9]), k, KITTI_FRONT_64x512))
    # Q6: density table 1..40 points in consecutive cells
    rows = []
    for c in range(1, 41):
        for j in range(c):
            rows.append([10.05 + 0.1 * c, 3.05, -2.0 + 0.01 * j, 0.5])
    cases.append(("q6_density", P(*rows), k, KITTI_FRONT_64x512))
    # Q8: front view last writer wins, values truncated to integers
    cases.append(("q8_front_last", P([20.0, 1.0, -1.0, 0.2], [20.0, 1.0, -1.0, 0.75], [20.0, 1.0, -1.00001, 0.4]),
                  k, KITTI_FRONT_64x512))
    # Q9: r_coord == R_SIZE (dropped: the reference stores out of bounds) and c_coord == C_SIZE (aliases next row)
    steep = np.tan(np.deg2rad(13.45)) * 1.0
    cases.append(("q9_front_bounds", P([1.0, 0.02, 0.79, 0.5], [3.0, 3.0, -1.0, 0.6], [3.0, 2.999, -1.0, 0.7],
                                       [3.0, -3.0, -1.0, 0.8], [3.0, 0.2, steep * 3.0, 0.9], [0.55, 0.001, 0.14, 0.3]),
                  k, KITTI_FRONT_64x512))
    # Q10: non-finite and huge coordinates are rejected (x86 float->int gives INT_MIN)
    nan, inf = np.nan, np.inf
    cases.append(("q10_nonfinite", P([nan, 0.0, -1.0, 0.5], [1.0, nan, -1.0, 0.5], [1.0, 1.0, nan, 0.5],
                                     [inf, 0.0, -1.0, 0.5], [1.0, -inf, -1.0, 0.5], [1.0, 1.0, inf, 0.5],
                                     [3e9, 1.0, -1.0, 0.5], [1.0, 3e9, -1.0, 0.5], [1.0, 1.0, -3e9, 0.5],
                                     [4.05, 4.05, -1.0, nan], [5.05, 4.05, -1.0, inf], [6.05, 4.05, -1.0, 3e9],
                                     [7.05, 4.05, -1.0, -0.5], [8.05, 4.05, -1.0, 0.5]), k, KITTI_FRONT_64x512))
    # one cell hammered by many points (atomic contention, count saturation at 31)
    rng = np.random.default_rng(7)
    hot = np.empty((5000, 4), np.float32)
    hot[:, 0] = 12.34 + rng.uniform(0, 0.09, 5000)
    hot[:, 1] = -7.77 + rng.uniform(0, 0.05, 5000)
    hot[:, 2] = rng.uniform(-4.2, 0.8, 5000)
    hot[:, 3] = rng.random(5000)
    hot[::7, 2] = np.float32(0.5)  # mass ties on the maximum
    cases.append(("hot_cell", hot, k, KITTI_FRONT_64x512))
    return cases


def small_cases():
    """Cheap cases on tiny grids: empty input, one point, dense random clouds with margins, ties."""
    cases = [
        ("empty", np.zeros((0, 4), np.float32), SMALL_TOP, SMALL_FRONT),
        ("single", P([1.0, 0.3, 0.1, 0.5]), SMALL_TOP, SMALL_FRONT),
        ("thin_grid", uniform_in_box(500, THIN_TOP, seed=3, margin=0.5), THIN_TOP, SMALL_FRONT),
        ("small_random", uniform_in_box(4000, SMALL_TOP, seed=1, margin=0.3), SMALL_TOP, SMALL_FRONT),
    ]
    q = uniform_in_box(3000, SMALL_TOP, seed=2, margin=0.1)
    q[:, 2] = np.round(q[:, 2], 1)  # mass ties on height
    cases.append(("small_ties", q, SMALL_TOP, SMALL_FRONT))
    return cases


def sweep_cases(frames=(0, 1)):
    cases = [(f"hdl64_{f}", hdl64_sweep(f), KITTI_PAPER_TOP, KITTI_FRONT_64x512) for f in frames]
    s0 = hdl64_sweep(0)
    perm = np.random.default_rng(42).permutation(s0.shape[0])
    cases.append(("hdl64_0_permuted", np.ascontiguousarray(s0[perm]), KITTI_PAPER_TOP, KITTI_FRONT_64x512))
    r = s0.copy()
    r[:, 2] = np.round(r[:, 2], 2)
    cases.append(("hdl64_0_zround", r, KITTI_PAPER_TOP, KITTI_FRONT_64x512))
    cases.append(("hdl64_0_refkitti", s0, REFERENCE_KITTI_TOP, KITTI_FRONT_64x512))
    cases.append(("hdl64_0_didi", s0, REFERENCE_DIDI_TOP, FrontParams(-20.0, 20.0, -45.0, 45.0, 0.1, 0.2)))
    cases.append(("uniform_in_box", uniform_in_box(120000, KITTI_PAPER_TOP, seed=5), KITTI_PAPER_TOP, KITTI_FRONT_64x512))
    return cases


def fuzz_cases(count=40, seed=2024):
    """Random parameter sets and clouds: odd plane counts (tiles that are not a multiple of 16 bytes), one-cell and
    no-slice grids, negative origins, coordinates sitting on cell and bin boundaries, ties, non-finite rows."""
    rng = np.random.default_rng(seed)
    divs = [0.05, 0.1, 0.2, 0.25, 0.3, 0.37, 0.5, 1.0]
    out = []
    for k in range(count):
        xn, yn, zn = int(rng.integers(1, 48)), int(rng.integers(1, 56)), int(rng.integers(0, 34))
        xd, yd, zd = (float(np.float32(rng.choice(divs))) for _ in range(3))
        x0, y0, z0 = (float(np.float32(rng.uniform(-20.0, 5.0))) for _ in range(3))
        top = TopParams(x0, x0 + xn * xd, y0, y0 + yn * yd, z0, z0 + max(zn, 1) * zd, xd, yd, zd, xn, yn, zn)
        rn, cn = int(rng.integers(1, 40)), int(rng.integers(1, 64))
        front = FrontParams(-20.0, 20.0, -45.0, 45.0, float(np.float32(rng.uniform(0.3, 6.0))),
                            float(np.float32(rng.uniform(0.3, 6.0))), rn, cn, 3)
        n = int(rng.choice([0, 1, 17, 300, 5000]))
        pts = np.empty((n, 4), np.float32)
        for a, (lo, d, cnt) in enumerate(((x0, xd, xn), (y0, yd, yn), (z0, zd, max(zn, 1)))):
            pts[:, a] = rng.uniform(lo - 2 * d, lo + (cnt + 2) * d, n)
        pts[:, 3] = rng.random(n)
        if n >= 17:
            snap = rng.random(n) < 0.3  # on cell boundaries: lo + j * div evaluated in float32
            for a, (lo, d) in enumerate(((x0, xd), (y0, yd), (z0, zd))):
                j = np.round((pts[snap, a] - np.float32(lo)) / np.float32(d))
                pts[snap, a] = np.float32(lo) + j.astype(np.float32) * np.float32(d)
            dup = rng.integers(0, n, n // 8)  # ties: copies of earlier points with another intensity
            pts[dup[: len(dup) // 2], :3] = pts[dup[len(dup) // 2: 2 * (len(dup) // 2)], :3]
            pts[rng.integers(0, n, 3), rng.integers(0, 4, 3)] = [np.nan, np.inf, -np.inf]
        out.append((f"fuzz{k}_{xn}x{yn}x{zn}_n{n}", pts, top, front))
    return out
