"""CPU tests of the empty-anchor oracle (oracle/anchors.py, oracle/anchor_oracle.c): the C restatement against the
numpy restatement, the clamping / 50-row / inverted-box rules of remove_empty_box_kernel.cu:16-40 as known answers,
numpy's channel sum restated, and the anchor generator against the reference's own `make_anchors`."""
import ast
import hashlib
import os

import numpy as np
import pytest

import anchor_cases

REFERENCE_RPN = "/root/reference/src/net/rpn_target_op.py"


@pytest.fixture(scope="module")
def anchors_oracle():
    from oracle import anchors
    anchors.build()
    return anchors


@pytest.mark.parametrize("shape", [(40, 30, 5), (64, 64, 1), (90, 70, 27), (33, 21, 40)])
def test_c_restatement_equals_numpy_restatement(anchors_oracle, shape):
    view = anchor_cases.sparse_view(shape, seed=1, fill=0.3, negative=True)
    boxes = anchor_cases.random_anchors(400, shape, seed=2)
    a = anchors_oracle.channel_sums(view, boxes)
    b = anchors_oracle.channel_sums_numpy(view, boxes)
    assert np.array_equal(a, b)


def test_known_answers(anchors_oracle):
    view = np.zeros((60, 20, 3), dtype=np.float32)
    view[10, 5] = (1, 2, 4)
    view[59, 19] = (8, 8, 8)          # the last row / column can never be inside a box: x2, y2 clamp to dim-1, exclusive
    view[12, 0] = (16, 0, 0)
    boxes = np.array([
        [5, 10, 6, 11],               # (y1, x1, y2, x2): exactly the cell (10, 5)
        [6, 10, 9, 12],               # next to it
        [0, 0, 100, 100],             # clamped to rows [0,59) x columns [0,19): but only the first 50 rows count
        [0, 11, 100, 100],            # rows 11..58: 48 rows, holds (12, 0) only
        [-7, -7, 6, 11],              # negative corner clamps to 0
        [6, 11, 5, 10],               # inverted: no rows
        [5, 10, 5, 11],               # zero columns
        [19, 59, 25, 65],             # entirely past the end: clamps to an empty box at the last cell
    ], dtype=np.int32)
    inds = anchors_oracle.channel_sums(view, boxes)
    expect = np.array([[1, 2, 4], [0, 0, 0], [17, 2, 4], [16, 0, 0], [1, 2, 4], [0, 0, 0], [0, 0, 0], [0, 0, 0]],
                      dtype=np.float32)
    assert np.array_equal(inds, expect)
    assert anchors_oracle.remove_empty_anchor(view, boxes, 0).tolist() == [0, 2, 3, 4]
    assert anchors_oracle.remove_empty_anchor(view, boxes, 7.0).tolist() == [2, 3]
    assert anchors_oracle.remove_empty_anchor(view, boxes, -1.0).tolist() == list(range(8))


def test_row_cap_is_50(anchors_oracle):
    view = np.ones((80, 4, 2), dtype=np.float32)
    boxes = np.array([[0, 0, 3, 49], [0, 0, 3, 50], [0, 0, 3, 51], [0, 5, 3, 79]], dtype=np.int32)
    inds = anchors_oracle.channel_sums(view, boxes)
    assert inds[:, 0].tolist() == [49 * 3, 50 * 3, 50 * 3, 50 * 3]


def test_summation_order_is_row_major_left_to_right(anchors_oracle):
    # 2^24 + 1 + 1 ... : a float32 running sum absorbs the ones, a sum that reached them first would not
    view = np.zeros((4, 8, 1), dtype=np.float32)
    view[0, 0, 0] = 2.0 ** 24
    view[0, 1:, 0] = 1.0
    view[1, :, 0] = 1.0
    box = np.array([[0, 0, 7, 3]], dtype=np.int32)
    got = anchors_oracle.channel_sums(view, box)[0, 0]
    row0 = np.float32(2.0 ** 24)            # + six ones, each absorbed
    row1 = np.float32(7.0)
    assert got == np.float32(np.float32(0) + row0) + row1 == np.float32(2.0 ** 24 + 8)


@pytest.mark.parametrize("c", [1, 2, 7, 8, 9, 15, 16, 17, 27, 31, 32, 33, 64, 127, 128, 129, 130, 257, 1000])
def test_pairwise_restatement_equals_numpy_sum(anchors_oracle, c):
    rng = np.random.default_rng(c)
    rows = (rng.standard_normal((60, c)) * rng.choice([1e-3, 1.0, 1e4], size=(60, c))).astype(np.float32)
    want = np.sum(rows, axis=1)
    got = np.array([anchors_oracle.pairwise_sum_f32(r) for r in rows], dtype=np.float32)
    assert np.array_equal(want, got)


def test_numpy_row_sum_does_not_depend_on_the_number_of_rows():
    rng = np.random.default_rng(5)
    rows = rng.standard_normal((120000, 27)).astype(np.float32)
    whole = np.sum(rows, axis=1)
    assert np.array_equal(whole[:100], np.sum(rows[:100].copy(), axis=1))
    assert np.array_equal(whole[-7:], np.sum(rows[-7:].copy(), axis=1))


def test_nonnegative_view_decision_is_occupancy(anchors_oracle):
    """The property the CUDA fast path rests on: without negative values, total > 0 <=> a positive cell in the box."""
    shape = (120, 90, 6)
    view = anchor_cases.sparse_view(shape, seed=3, fill=0.02)
    view[view > 0] *= np.float32(1e-30)      # tiny values must still count
    boxes = anchor_cases.random_anchors(3000, shape, seed=4)
    kept = set(anchors_oracle.remove_empty_anchor(view, boxes, 0.0).tolist())
    occupied = view.max(axis=2) > 0
    for a, box in enumerate(boxes):
        x1, x2, y1, y2 = anchors_oracle.clamp_box(box, shape[0], shape[1])
        x2 = min(x2, x1 + anchors_oracle.MAX_LINES)
        assert (a in kept) == bool(occupied[x1:x2, y1:y2].any()), a


def test_make_anchors_equals_reference():
    from mv3d_b200.anchors import MV3D_BASES, make_anchors
    mine = make_anchors((800, 600), stride=8)
    assert mine.shape == (100 * 75 * 4, 4) and mine.dtype == np.int32
    assert hashlib.sha256(mine.tobytes()).hexdigest()[:16] == "0f4011d169a318f0"
    if not os.path.exists(REFERENCE_RPN):
        return
    tree = ast.parse(open(REFERENCE_RPN).read())
    fn = next(n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "make_anchors")
    ns = {"np": np}
    exec(compile(ast.Module(body=[fn], type_ignores=[]), REFERENCE_RPN, "exec"), ns)
    theirs, _ = ns["make_anchors"](MV3D_BASES, 8, (800, 600), (100, 75))   # src/mv3d.py:230
    assert np.array_equal(mine, theirs)
