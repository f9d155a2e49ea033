"""Error behaviour on a machine WITH a GPU ("Errors"): the reference routines return void and validate
nothing; the replacement never throws or aborts across the ABI -- a bad call returns non-zero (Part 2) or leaves the
outputs zero-filled (the reference symbols), records the reason for mv3d_last_error(), and the next good call works."""
import ctypes

import numpy as np
import pytest

from mv3d_b200.params import KITTI_FRONT_64x512, KITTI_PAPER_TOP, FrontParams, TopParams
from mv3d_b200.sweeps import hdl64_sweep, pack_sweeps

pytestmark = pytest.mark.gpu
SMALL = TopParams(0.0, 4.0, -2.4, 2.4, -1.0, 1.0, 0.2, 0.2, 0.4, 20, 24, 5)
SMALL_FRONT = FrontParams(-20.0, 20.0, -45.0, 45.0, 2.5, 5.0, 16, 18, 3)


def _good_call(enc, torch, oracle_mod):
    s = hdl64_sweep(5)[:5000]
    pts, offs = pack_sweeps([s])
    top, front = enc.encode_device(torch.from_numpy(pts).cuda(), offs, KITTI_PAPER_TOP, KITTI_FRONT_64x512)
    torch.cuda.synchronize()
    et, ef = oracle_mod.top_and_front_maps(s, KITTI_PAPER_TOP, KITTI_FRONT_64x512)
    assert np.array_equal(top[0].cpu().numpy(), et) and np.array_equal(front[0].cpu().numpy(), ef)


def test_bad_device_calls_fail_and_the_context_survives(oracle_mod):
    import torch
    from mv3d_b200.encoder import ViewEncoder
    enc = ViewEncoder(0)
    d = torch.zeros((100, 4), device="cuda")
    with pytest.raises(RuntimeError, match="offsets decrease"):
        enc.encode_device(d, np.array([0, 60, 40], np.int32), SMALL, SMALL_FRONT)
    with pytest.raises(RuntimeError, match="negative"):
        enc.encode_device(d, np.array([-1, 40], np.int32), SMALL, SMALL_FRONT)
    with pytest.raises(RuntimeError, match="integral"):  # F_SIZE < 3: no room for the three channels
        enc.encode_device(d, np.array([0, 100], np.int32), SMALL, FrontParams(-20.0, 20.0, -45.0, 45.0, 2.5, 5.0, 16, 18, 2))
    # nothing to do is not an error
    t, f = enc.encode_device(d[:0], np.array([0], np.int32), SMALL, SMALL_FRONT)
    assert t.shape[0] == 0 and f.shape[0] == 0
    _good_call(enc, torch, oracle_mod)
    enc.close()


def test_raw_abi_rejects_null_and_unknown_arguments(oracle_mod):
    import torch
    from mv3d_b200 import _lib
    from mv3d_b200.encoder import ViewEncoder, _front_c, _top_c
    lib = _lib.load_library()
    enc = ViewEncoder(0)
    ctx = enc._ctx
    offs = (ctypes.c_int * 2)(0, 10)
    d = torch.zeros((10, 4), device="cuda")
    top = torch.zeros((1, *SMALL.shape), device="cuda")
    tp, fp = _top_c(SMALL), _front_c(SMALL_FRONT)
    call = lib.mv3d_encode_batch_dev
    assert call(None, d.data_ptr(), offs, 1, top.data_ptr(), None, ctypes.byref(tp), None, 0, 0, None) != 0
    assert "ctx is NULL" in _lib.last_error()
    assert call(ctx, d.data_ptr(), None, 1, top.data_ptr(), None, ctypes.byref(tp), None, 0, 0, None) != 0
    assert "offsets is NULL" in _lib.last_error()
    assert call(ctx, d.data_ptr(), offs, -1, top.data_ptr(), None, ctypes.byref(tp), None, 0, 0, None) != 0
    assert "negative batch" in _lib.last_error()
    assert call(ctx, d.data_ptr(), offs, 1, top.data_ptr(), None, None, None, 0, 0, None) != 0
    assert "top parameters are required" in _lib.last_error()
    assert call(ctx, d.data_ptr(), offs, 1, top.data_ptr(), None, ctypes.byref(tp), None, 7, 0, None) != 0
    assert "unknown semantics" in _lib.last_error()
    bad = _top_c(SMALL)
    bad.yn = -3
    assert call(ctx, d.data_ptr(), offs, 1, top.data_ptr(), None, ctypes.byref(bad), None, 0, 0, None) != 0
    assert "negative grid" in _lib.last_error()
    front = torch.zeros((1, *SMALL_FRONT.shape), device="cuda")
    assert call(ctx, d.data_ptr(), offs, 1, top.data_ptr(), front.data_ptr(), ctypes.byref(tp), None, 1, 0, None) != 0
    assert "front parameters are required" in _lib.last_error()
    assert call(ctx, d.data_ptr(), offs, 1, top.data_ptr(), front.data_ptr(), ctypes.byref(tp), ctypes.byref(fp), 1, 0, None) == 0
    torch.cuda.synchronize()
    _good_call(enc, torch, oracle_mod)
    enc.close()


def test_reference_symbol_with_bad_sizes_zero_fills(oracle_mod):
    """`C_SIZE, R_SIZE, F_SIZE` travel as floats (LidarTopAndFrontPreprocess.c:30); a non-integral one cannot index an
    image. The symbol returns void: the outputs come back zero-filled and the reason is on record."""
    from mv3d_b200 import _lib
    from mv3d_b200.ver3 import createTopAndFrontMaps
    s = hdl64_sweep(6)[:2000]
    top = np.full(SMALL.shape, 9.0, np.float32)
    front = np.full((16, 18, 3), 9.0, np.float32)
    lib = _lib.load_library()
    F, I = ctypes.c_float, ctypes.c_int
    t = SMALL
    lib.createTopAndFrontMaps(ctypes.c_void_p(s.ctypes.data), I(s.shape[0]), ctypes.c_void_p(top.ctypes.data),
                              ctypes.c_void_p(front.ctypes.data), F(t.x_min), F(t.x_max), F(t.y_min), F(t.y_max), F(t.z_min),
                              F(t.z_max), F(t.x_div), F(t.y_div), F(t.z_div), I(t.xn), I(t.yn), I(t.zn), F(-20.0), F(20.0),
                              F(-45.0), F(45.0), F(2.5), F(5.0), F(18.5), F(16.0), F(3.0))  # C_SIZE = 18.5
    assert "integral" in _lib.last_error()
    assert not top.any() and not front.any()
    createTopAndFrontMaps(s, s.shape[0], top, front, SMALL.as_top_paras(), SMALL_FRONT.as_front_paras())
    et, ef = oracle_mod.top_and_front_maps(s, SMALL, SMALL_FRONT)
    assert np.array_equal(top, et) and np.array_equal(front, ef)
    assert _lib.last_error() == ""


@pytest.mark.parametrize("xn,yn", [(0, 24), (20, 0), (0, 0)])
def test_empty_grid_is_a_no_op(oracle_mod, xn, yn):
    """X_SIZE = 0 or Y_SIZE = 0: the reference's loops simply do not run (LidarTopPreprocess.c:59-84,112). Here the tile
    count of an empty grid must not divide by zero (round-1 advisor finding); the front view of such a call is all
    zeros too, because nothing passes the bird's-eye-view gate."""
    import torch
    from mv3d_b200.encoder import ViewEncoder
    from mv3d_b200.pytwin import PyTwinEncoder
    from mv3d_b200.params import PYTWIN_KITTI, replace
    from mv3d_b200.ver3 import createTopAndFrontMaps, createTopMaps
    top = TopParams(0.0, 4.0, -2.4, 2.4, -1.0, 1.0, 0.2, 0.2, 0.4, xn, yn, 5)
    s = hdl64_sweep(3)[:3000]
    pts, offs = pack_sweeps([s, s[:10]])
    d = torch.from_numpy(pts).cuda()
    for strategy in ("auto", "tile", "atomic", "sort", "patch"):
        enc = ViewEncoder(0, strategy=strategy)
        t, f = enc.encode_device(d, offs, top, SMALL_FRONT)
        torch.cuda.synchronize()
        assert t.numel() == 0 and tuple(t.shape) == (2, xn, yn, 7)
        assert not f.any().item(), strategy
        enc.close()
    out = np.ones(top.shape, np.float32)
    fo = np.ones(SMALL_FRONT.shape, np.float32)
    createTopMaps(s, s.shape[0], out, top.as_top_paras())
    createTopAndFrontMaps(s, s.shape[0], out, fo, top.as_top_paras(), SMALL_FRONT.as_front_paras())
    et, ef = oracle_mod.top_and_front_maps(s, top, SMALL_FRONT)
    assert np.array_equal(fo, ef) and not fo.any()
    py = PyTwinEncoder(0)  # the Python-twin parameters derive their grid from the ranges: (-0.05 // 0.1) + 1 = 0 rows
    empty = replace(PYTWIN_KITTI, x_max=PYTWIN_KITTI.x_min - 0.05)
    assert empty.xn == 0
    ptop, pfront = py.encode_device(d, offs, empty)
    torch.cuda.synchronize()
    assert ptop.numel() == 0 and not pfront.any().item()  # nothing survives the crop 0 < x < -0.05 either
    py.close()


def test_wrapper_refuses_mismatched_buffers():
    """The Python mirror checks what the C ABI cannot: offsets against the point buffer, output shape / dtype / device.
    A mismatch is an exception here instead of an out-of-bounds access on the device (round-1 advisor finding)."""
    import torch
    from mv3d_b200.encoder import ViewEncoder
    enc = ViewEncoder(0)
    d = torch.zeros((100, 4), device="cuda")
    with pytest.raises(ValueError, match="point buffer holds"):
        enc.encode_device(d, np.array([0, 101], np.int32), SMALL, SMALL_FRONT)
    with pytest.raises(ValueError, match="shape"):
        enc.encode_device(d, np.array([0, 100], np.int32), SMALL, SMALL_FRONT, top_out=torch.zeros((1, 20, 24, 6), device="cuda"))
    with pytest.raises(ValueError, match="float32 CUDA tensor"):
        enc.encode_device(d, np.array([0, 100], np.int32), SMALL, SMALL_FRONT, top_out=torch.zeros((1, *SMALL.shape)))
    with pytest.raises(ValueError, match="float32 CUDA tensor"):
        enc.encode_device(d, np.array([0, 100], np.int32), SMALL, SMALL_FRONT,
                          front_out=torch.zeros((1, *SMALL_FRONT.shape), device="cuda", dtype=torch.float64))
    h = np.zeros((100, 4), np.float32)
    with pytest.raises(ValueError, match="point buffer holds"):
        enc.encode_host(h, np.array([0, 400], np.int32), SMALL, SMALL_FRONT)
    with pytest.raises(ValueError, match="shape"):
        enc.encode_host(h, np.array([0, 100], np.int32), SMALL, SMALL_FRONT, top_out=np.zeros((2, *SMALL.shape), np.float32))
    with pytest.raises(ValueError, match="float32 numpy"):
        enc.encode_host(h, np.array([0, 100], np.int32), SMALL, SMALL_FRONT, top_out=np.zeros((1, *SMALL.shape), np.float64))
    enc.close()


def test_selfcheck_reports_and_clears(torch_cuda, encoder_lib):
    """The library's own bounds checks (csrc/selfcheck.cuh, -DMV3D_SELF_CHECK: the stand-in for compute-sanitizer on this
    pool): on such a build a deliberate violation is reported once and cleared; on a normal build the hook says so."""
    import ctypes
    lib = encoder_lib
    ctx = ctypes.c_void_p()
    assert lib.mv3d_ctx_create(0, ctypes.byref(ctx)) == 0
    flags = ctypes.c_uint(123)
    enabled = lib.mv3d_debug_selfcheck(ctx, ctypes.byref(flags))
    assert enabled in (0, 1) and flags.value == 0
    assert lib.mv3d_debug_selfcheck_trip(ctx) == 0
    assert lib.mv3d_debug_selfcheck(ctx, ctypes.byref(flags)) == enabled
    assert flags.value == ((1 << 15) if enabled else 0)
    assert lib.mv3d_debug_selfcheck(ctx, ctypes.byref(flags)) == enabled and flags.value == 0
    lib.mv3d_ctx_destroy(ctx)
