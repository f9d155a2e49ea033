"""GPU parity tests of the fused caller-side steps of `clidar_to_top` (src/data.py:265-291): the centre-car filter
(:224-227) before the C routine and np.flipud(np.fliplr(.)) after it, against the oracle wrapped in exactly those
numpy steps."""
import numpy as np
import pytest

from mv3d_b200.params import (CLIDAR_DIDI, CLIDAR_DIDI2, CLIDAR_KITTI, CLIDAR_PRESETS, KITTI_FRONT_64x512,
                              REFERENCE_DIDI_TOP, replace)
from mv3d_b200.sweeps import hdl64_sweep, pack_sweeps

pytestmark = pytest.mark.gpu


def filter_center_car(lidar):
    """src/data.py:224-227, as numpy evaluates it on a float32 array."""
    keep = np.logical_or(np.abs(lidar[:, 0]) > 4.7 / 2, np.abs(lidar[:, 1]) > 2.1 / 2)
    return lidar[keep]


def sweep_with_car_points(frame, seed):
    """An HDL-64 sweep plus points on and around the ego vehicle's footprint (incl. exactly on its edges)."""
    rng = np.random.default_rng(seed)
    near = rng.uniform(-4.0, 4.0, size=(4000, 4)).astype(np.float32)
    near[:, 2] = rng.uniform(-2.5, 0.5, size=4000)
    near[:, 3] = rng.random(4000)
    edge = np.array([[2.35, 0.0, -1.0, 0.5], [-2.35, 0.2, -1.0, 0.5], [0.1, 1.05, -1.0, 0.5], [0.1, -1.05, -1.0, 0.5],
                     [np.float32(2.35), 1.0, -1.0, 0.7], [np.nextafter(np.float32(2.35), np.float32(3)), 0.0, -1.0, 0.9]],
                    dtype=np.float32)
    return np.concatenate([hdl64_sweep(frame)[::3], near, edge]).astype(np.float32)


def expected_top(oracle_mod, pts, top):
    raw = filter_center_car(pts) if top.filter_center_car else pts
    plain = replace(top, flip=False, filter_center_car=False)
    out = oracle_mod.top_maps(np.ascontiguousarray(raw), plain)
    return np.ascontiguousarray(out[::-1, ::-1]) if top.flip else out


@pytest.mark.parametrize("preset", ["kitti", "didi", "didi2"])
@pytest.mark.parametrize("strategy", ["tile", "atomic"])
def test_clidar_presets_device(oracle_mod, preset, strategy):
    import torch
    from mv3d_b200.encoder import ViewEncoder
    top = CLIDAR_PRESETS[preset]
    sweeps = [sweep_with_car_points(f, 10 + f) for f in range(3)] + [np.zeros((0, 4), np.float32)]
    pts, offs = pack_sweeps(sweeps)
    enc = ViewEncoder(0, semantics="top", strategy=strategy)
    out, _ = enc.encode_device(torch.from_numpy(pts).cuda(), offs, top)
    torch.cuda.synchronize()
    out = out.cpu().numpy()
    for f, s in enumerate(sweeps):
        want = expected_top(oracle_mod, s, top)
        assert np.array_equal(out[f], want), (preset, strategy, f)
    if top.filter_center_car:  # the filter really removed something that the box would have kept
        assert not np.array_equal(expected_top(oracle_mod, sweeps[0], replace(top, filter_center_car=False)), out[0])
    enc.close()


@pytest.mark.parametrize("flags", [(True, False), (False, True), (True, True)])
def test_flags_with_front_view(oracle_mod, flags):
    """createTopAndFrontMaps arithmetic: the front view sees only the points that survive the filter (it is computed
    from the array the caller passes) and is not flipped."""
    import torch
    from mv3d_b200.encoder import ViewEncoder
    flip, center = flags
    top = replace(REFERENCE_DIDI_TOP, flip=flip, filter_center_car=center)
    sweeps = [sweep_with_car_points(f, 20 + f) for f in range(2)]
    pts, offs = pack_sweeps(sweeps)
    enc = ViewEncoder(0, semantics="topfront", strategy="tile")
    out, fout = enc.encode_device(torch.from_numpy(pts).cuda(), offs, top, KITTI_FRONT_64x512)
    torch.cuda.synchronize()
    for f, s in enumerate(sweeps):
        raw = np.ascontiguousarray(filter_center_car(s) if center else s)
        et, ef = oracle_mod.top_and_front_maps(raw, replace(top, flip=False, filter_center_car=False), KITTI_FRONT_64x512)
        if flip:
            et = np.ascontiguousarray(et[::-1, ::-1])
        assert np.array_equal(out[f].cpu().numpy(), et), f
        assert np.array_equal(fout[f].cpu().numpy(), ef), f
    enc.close()


@pytest.mark.parametrize("readback", ["compact", "dense"])
def test_clidar_to_top_host(oracle_mod, readback, monkeypatch):
    """The reference-named wrapper, numpy in / numpy out, for every data-set preset."""
    import mv3d_b200.ver3 as ver3
    monkeypatch.setenv("MV3D_HOST_READBACK", readback)
    monkeypatch.setattr(ver3, "_clidar_encoder", None)
    s = sweep_with_car_points(5, 30)
    for top in (CLIDAR_KITTI, CLIDAR_DIDI, CLIDAR_DIDI2):
        got = ver3.clidar_to_top(s, top)
        assert got.dtype == np.float32 and got.flags["C_CONTIGUOUS"]
        assert np.array_equal(got, expected_top(oracle_mod, s, top))
    monkeypatch.setattr(ver3, "_clidar_encoder", None)


def test_unknown_flags_are_rejected():
    import ctypes
    import torch
    from mv3d_b200 import _lib
    from mv3d_b200.encoder import ViewEncoder, _top_c
    enc = ViewEncoder(0, semantics="top")
    tc = _top_c(CLIDAR_KITTI)
    tc.flags = 8
    pts = torch.zeros((4, 4), device="cuda")
    out = torch.zeros((1, *CLIDAR_KITTI.shape), device="cuda")
    offs = (ctypes.c_int * 2)(0, 4)
    rc = enc._lib.mv3d_encode_batch_dev(enc._ctx, pts.data_ptr(), offs, 1, out.data_ptr(), None, ctypes.byref(tc), None,
                                        _lib.SEM_TOP, _lib.STRATEGY_AUTO, None)
    assert rc != 0 and "flags" in _lib.last_error()
    enc.close()
