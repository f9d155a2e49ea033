"""GPU parity tests of the empty-anchor filter (include/mv3d_anchors.h) against the oracle and against the
reference's own kernel executed on this GPU (oracle/_ref/remove_empty_box_kernel.cubin)."""
import numpy as np
import pytest

import anchor_cases

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def oracle_anchors():
    from oracle import anchors
    anchors.build()
    return anchors


@pytest.fixture(scope="module")
def filt():
    from mv3d_b200.anchors import AnchorFilter
    f = AnchorFilter(0)
    yield f
    f.close()


def _run(filt, views, boxes, limit=0.0, mode="auto", want_totals=False):
    import torch
    from mv3d_b200 import _lib
    filt.mode = {"auto": _lib.ANCHORS_AUTO, "exact": _lib.ANCHORS_EXACT}[mode]
    filt.set_anchors(boxes)
    v = torch.from_numpy(np.ascontiguousarray(views)).cuda()
    out = filt.filter_device(v, limit, want_totals=want_totals)
    torch.cuda.synchronize()
    keep, count = out[0].cpu().numpy(), out[1].cpu().numpy()
    kept = [keep[f, : count[f]] for f in range(keep.shape[0])]
    return (kept, out[2].cpu().numpy()) if want_totals else kept


# -- reference execution ------------------------------------------------------------------------------------
def test_reference_kernel_one_row_boxes(oracle_anchors):
    """Boxes one row high: a single thread block per anchor adds to inds[], no race -> the executed reference
    kernel must equal the oracle bit for bit (row integration order, clamping, channel layout)."""
    import ref_anchor_kernel as ref
    if not ref.available():
        pytest.skip("oracle/_ref/remove_empty_box_kernel.cubin or cuda-python not present")
    shape = (90, 130, 27)
    view = anchor_cases.dense_view(shape, seed=11)
    boxes = anchor_cases.line_anchors(5000, shape, seed=12)
    assert np.array_equal(ref.channel_sums(view, boxes), oracle_anchors.channel_sums(view, boxes))


def test_reference_kernel_tall_boxes_integer_view(oracle_anchors):
    """Taller boxes race in the reference (`+=` from up to 50 blocks). On a non-negative integer-valued view
    every schedule without a lost update gives the oracle's sum exactly and a lost update can only lower it."""
    import ref_anchor_kernel as ref
    if not ref.available():
        pytest.skip("oracle/_ref/remove_empty_box_kernel.cubin or cuda-python not present")
    shape = (120, 100, 8)
    view = anchor_cases.sparse_view(shape, seed=13, fill=0.2, integer=True)
    boxes = anchor_cases.random_anchors(4000, shape, seed=14)
    theirs, ours = ref.channel_sums(view, boxes), oracle_anchors.channel_sums(view, boxes)
    assert np.all(theirs <= ours)
    same = float(np.mean(np.all(theirs == ours, axis=1)))
    print(f"reference kernel == oracle on {100 * same:.2f} % of the anchors (the rest lost an update to its race)")
    empty = ~np.any(ours > 0, axis=1)
    assert np.all(theirs[empty] == 0)          # an empty box is empty under every schedule
    assert same > 0.5


# -- device path vs oracle ----------------------------------------------------------------------------------
@pytest.mark.parametrize("shape", [(800, 600, 27), (97, 61, 27), (64, 64, 1), (50, 37, 3), (40, 40, 33)])
def test_auto_on_nonnegative_views(filt, oracle_anchors, shape):
    views = np.stack([anchor_cases.sparse_view(shape, seed=20 + f, fill=0.01 * (f + 1)) for f in range(3)])
    boxes = anchor_cases.random_anchors(20000 if shape[0] > 100 else 3000, shape, seed=21)
    kept = _run(filt, views, boxes)
    for f in range(3):
        assert np.array_equal(kept[f], oracle_anchors.remove_empty_anchor(views[f], boxes, 0.0)), f


def test_auto_counts_tiny_values(filt, oracle_anchors):
    shape = (100, 80, 5)
    view = anchor_cases.sparse_view(shape, seed=30, fill=0.01)
    view[view > 0] = np.float32(1e-44)        # denormals are positive
    boxes = anchor_cases.random_anchors(5000, shape, seed=31)
    want = oracle_anchors.remove_empty_anchor(view, boxes, 0.0)
    assert len(want) > 0
    assert np.array_equal(_run(filt, view[None], boxes)[0], want)


@pytest.mark.parametrize("poison", ["negative", "minus_zero", "nan", "none"])
def test_auto_falls_back_per_frame(filt, oracle_anchors, poison):
    """A frame with a negative value, a -0 or a NaN takes the exact sums; its neighbours in the batch do not."""
    shape = (120, 90, 27)
    views = np.stack([anchor_cases.sparse_view(shape, seed=40 + f, fill=0.02) for f in range(3)])
    if poison == "negative":
        views[1] = anchor_cases.sparse_view(shape, seed=44, fill=0.05, negative=True)
    elif poison == "minus_zero":
        views[1, 7, 9, 3] = np.float32(-0.0)
    elif poison == "nan":
        views[1, 60, 45, 26] = np.nan
    boxes = anchor_cases.random_anchors(6000, shape, seed=45)
    kept = _run(filt, views, boxes)
    for f in range(3):
        with np.errstate(invalid="ignore"):
            want = oracle_anchors.remove_empty_anchor(views[f], boxes, 0.0)
        assert np.array_equal(kept[f], want), f


@pytest.mark.parametrize("shape,limit", [((800, 600, 27), 0.0), ((90, 70, 27), 0.01), ((90, 70, 27), -0.02),
                                         ((64, 48, 1), 0.0), ((50, 37, 7), 0.0), ((30, 30, 130), 0.05),
                                         ((20, 20, 300), 0.0)])
def test_exact_totals_bit_equal(filt, oracle_anchors, shape, limit):
    """The reference's own smoke input, 0.01*randn: every total must carry the reference's summation order."""
    view = anchor_cases.dense_view(shape, seed=50)
    boxes = anchor_cases.random_anchors(12000 if shape[0] > 100 else 3000, shape, seed=51)
    kept, totals = _run(filt, view[None], boxes, limit=limit, mode="exact", want_totals=True)
    want = oracle_anchors.anchor_totals(view, boxes)
    assert np.array_equal(totals[0], want)
    assert np.array_equal(kept[0], np.where(want > np.float32(limit))[0])


def test_auto_with_nonzero_limit_takes_sums(filt, oracle_anchors):
    shape = (100, 100, 27)
    view = anchor_cases.sparse_view(shape, seed=60, fill=0.05)
    boxes = anchor_cases.random_anchors(5000, shape, seed=61)
    for limit in (0.5, 30.0, -1.0):
        assert np.array_equal(_run(filt, view[None], boxes, limit=limit)[0],
                              oracle_anchors.remove_empty_anchor(view, boxes, limit)), limit


def test_reference_smoke_configuration(filt, oracle_anchors):
    """remove_empty_box.py:49-70: 120 000 boxes (0,0,40,16)+shift over 0.01*randn(800,600,27), limit 0."""
    view = anchor_cases.dense_view((800, 600, 27), seed=70)
    boxes = anchor_cases.smoke_anchors(120000, seed=71)
    from mv3d_b200.anchors import remove_empty_anchor
    got = remove_empty_anchor(view, boxes, 0)
    assert got.dtype == np.int64
    assert np.array_equal(got, oracle_anchors.remove_empty_anchor(view, boxes, 0))


def test_edge_cases(filt, oracle_anchors):
    import torch
    shape = (33, 21, 5)
    view = anchor_cases.sparse_view(shape, seed=80, fill=0.1)
    # no anchors at all
    filt.set_anchors(np.zeros((0, 4), np.int32))
    keep, count = filt.filter_device(torch.from_numpy(view).cuda())
    assert keep.shape == (1, 0) and int(count[0]) == 0
    # every anchor outside / inverted: nothing kept
    boxes = np.array([[100, 100, 200, 200], [5, 5, 2, 2], [-50, -50, -10, -10]], dtype=np.int32)
    assert _run(filt, view[None], boxes)[0].tolist() == []
    # everything kept, chunk boundary (1024) crossed, last chunk ragged
    full = np.ones(shape, dtype=np.float32)
    boxes = np.tile(np.array([[0, 0, 10, 10]], np.int32), (2500, 1))
    assert np.array_equal(_run(filt, full[None], boxes)[0], np.arange(2500))
    # a view that is not 16-byte aligned per frame (cells*planes % 4 != 0) in a batch
    odd = (7, 9, 3)
    views = np.stack([anchor_cases.sparse_view(odd, seed=81 + f, fill=0.2) for f in range(4)])
    boxes = anchor_cases.random_anchors(500, odd, seed=85, max_w=8, max_l=8, margin=3)
    kept = _run(filt, views, boxes)
    for f in range(4):
        assert np.array_equal(kept[f], oracle_anchors.remove_empty_anchor(views[f], boxes, 0.0))


def test_filters_the_encoders_output_in_place(filt, oracle_anchors, oracle_mod):
    """The path the row is about: encode on the device, filter the resident tensor, nothing leaves the GPU between."""
    import torch
    from mv3d_b200.anchors import AnchorFilter, make_anchors
    from mv3d_b200.encoder import ViewEncoder
    from mv3d_b200.params import REFERENCE_KITTI_TOP as top_p
    from mv3d_b200.sweeps import pack_sweeps, uniform_in_box
    sweeps = [uniform_in_box(20000 + 7000 * i, top_p, seed=90 + i, margin=2.0) for i in range(3)]
    pts, offs = pack_sweeps(sweeps)
    enc = ViewEncoder(0, semantics="top")
    shared = AnchorFilter(enc, anchors=make_anchors(top_p.shape[:2], stride=8))
    top, _ = enc.encode_device(torch.from_numpy(pts).cuda(), offs, top_p)
    keep, count = shared.filter_device(top)
    torch.cuda.synchronize()
    boxes = make_anchors(top_p.shape[:2], stride=8)
    assert boxes.shape[0] == 30000
    for f, s in enumerate(sweeps):
        want = oracle_anchors.remove_empty_anchor(oracle_mod.top_maps(s, top_p), boxes, 0.0)
        assert 0 < len(want) < 30000
        assert np.array_equal(keep[f, : int(count[f])].cpu().numpy(), want), f
    enc.close()


@pytest.mark.parametrize("strategy", ["tile", "atomic"])
@pytest.mark.parametrize("with_front", [False, True])
def test_fused_encode_and_filter_equals_two_calls(oracle_anchors, oracle_mod, strategy, with_front):
    """One call (the tile kernel reports occupancy while it builds the view) == encode, then filter == the oracle.
    70 sweeps: more than one launch chunk, so the per-chunk offsets into the occupancy map are exercised."""
    import torch
    from mv3d_b200.anchors import AnchorFilter, make_anchors
    from mv3d_b200.encoder import ViewEncoder
    from mv3d_b200.params import KITTI_FRONT_64x512 as front_p, REFERENCE_KITTI_TOP as top_p
    from mv3d_b200.sweeps import pack_sweeps, uniform_in_box
    sweeps = [uniform_in_box(3000 + 900 * i, top_p, seed=100 + i, margin=2.0) for i in range(70)]
    sweeps[5] = sweeps[5][:0]                                  # an empty sweep keeps nothing
    sweeps[9] = sweeps[9].copy()
    sweeps[9][:, 3] -= np.float32(0.5)                         # negative intensities: this frame needs the sums
    pts, offs = pack_sweeps(sweeps)
    boxes = make_anchors(top_p.shape[:2], stride=8)
    enc = ViewEncoder(0, semantics="topfront", strategy=strategy)
    filt = AnchorFilter(enc, anchors=boxes)
    d = torch.from_numpy(pts).cuda()
    fp = front_p if with_front else None
    top, front, keep, count = filt.encode_and_filter(d, offs, top_p, front=fp)
    top2, front2 = enc.encode_device(d, offs, top_p, fp)
    keep2, count2 = filt.filter_device(top2)
    torch.cuda.synchronize()
    assert torch.equal(top, top2) and torch.equal(count, count2)
    if with_front:
        assert torch.equal(front, front2)
    count = count.cpu().numpy()
    assert count[5] == 0
    for f in range(70):
        assert torch.equal(keep[f, : count[f]], keep2[f, : count[f]]), f
    for f in (0, 9, 69):
        et, _ = oracle_mod.top_and_front_maps(sweeps[f], top_p, front_p)
        want = oracle_anchors.remove_empty_anchor(et, boxes, 0.0)
        assert np.array_equal(keep[f, : count[f]].cpu().numpy(), want), f
    enc.close()


def test_fused_python_twin(oracle_anchors):
    import torch
    from oracle import pytwin
    from mv3d_b200.anchors import AnchorFilter, make_anchors
    from mv3d_b200.params import PYTWIN_KITTI as P
    from mv3d_b200.pytwin import PyTwinEncoder
    from mv3d_b200.sweeps import hdl64_sweep, pack_sweeps
    sweeps = [hdl64_sweep(f)[::7] for f in range(3)]
    pts, offs = pack_sweeps(sweeps)
    boxes = make_anchors(P.top_shape[:2], stride=8)
    enc = PyTwinEncoder(0)
    filt = AnchorFilter(enc, anchors=boxes)
    top, front, keep, count = filt.encode_and_filter(torch.from_numpy(pts).cuda(), offs, P)
    torch.cuda.synchronize()
    for f, s in enumerate(sweeps):
        et = pytwin.lidar_to_top(s, P)
        assert np.array_equal(top[f].cpu().numpy(), et)
        want = oracle_anchors.remove_empty_anchor(et, boxes, 0.0)
        assert np.array_equal(keep[f, : int(count[f])].cpu().numpy(), want), f
    enc.close()
