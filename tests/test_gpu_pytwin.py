"""GPU parity tests of the Python-twin mode: the CUDA path, through the C ABI, against
the numpy restatement of the reference's Python encoders (oracle/pytwin.py) and against the committed outputs of the
reference functions themselves (tests/golden/pytwin_golden.npz). Bit-exact on every channel."""
import os

import numpy as np
import pytest

import pytwin_cases
from golden.make_golden import unpack
from mv3d_b200.params import PYTWIN_DIDI2, PYTWIN_KITTI
from mv3d_b200.sweeps import hdl64_sweep, pack_sweeps

pytestmark = pytest.mark.gpu
CASES = pytwin_cases.all_cases()
GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "pytwin_golden.npz")


def same(a, b, what):
    assert a.shape == b.shape, what
    bad = ~((a == b) | (np.isnan(a) & np.isnan(b)))
    assert not bad.any(), f"{what}: {np.count_nonzero(bad)} of {a.size} differ, first at {np.argwhere(bad)[:3].tolist()}"


@pytest.fixture(scope="module")
def pytwin():
    from oracle import pytwin
    pytwin.build()
    return pytwin


@pytest.mark.parametrize("name,pts,p", CASES, ids=[c[0] for c in CASES])
def test_host_functions_vs_oracle_and_golden(pytwin, name, pts, p):
    """`lidar_to_top` / `lidar_to_front_fast` with the reference's names and conventions (numpy in, numpy out)."""
    from mv3d_b200.pytwin import lidar_to_front_fast, lidar_to_top
    g = np.load(GOLDEN)
    top, front = lidar_to_top(pts, p), lidar_to_front_fast(pts, p)
    same(top, pytwin.lidar_to_top(pts, p), f"{name} top vs oracle")
    same(front, pytwin.lidar_to_front_fast(pts, p), f"{name} front vs oracle")
    same(top, unpack(g, name, "top", p.top_shape), f"{name} top vs the reference's output")
    same(front, unpack(g, name, "front", p.front_shape), f"{name} front vs the reference's output")


def test_device_batch_ragged(pytwin):
    """A ragged batch in one launch, twice (scratch must clean itself), both presets."""
    import torch
    from mv3d_b200.pytwin import PyTwinEncoder
    enc = PyTwinEncoder(0)
    qc = {c[0]: c[1] for c in pytwin_cases.quirk_cases()}
    sweeps = [hdl64_sweep(3), np.zeros((0, 4), np.float32), qc["integral_qz"], hdl64_sweep(4)[:40001], qc["front_pixels"],
              qc["ties_density"], hdl64_sweep(5)[:1]]
    pts, offs = pack_sweeps(sweeps)
    d = torch.from_numpy(pts).cuda()
    for p in (PYTWIN_KITTI, PYTWIN_DIDI2):
        for rep in range(2):
            top, front = enc.encode_device(d, offs, p)
            torch.cuda.synchronize()
            top, front = top.cpu().numpy(), front.cpu().numpy()
            for i, s in enumerate(sweeps):
                same(top[i], pytwin.lidar_to_top(s, p), f"rep{rep} sweep {i} top")
                same(front[i], pytwin.lidar_to_front_fast(s, p), f"rep{rep} sweep {i} front")
    t, f = enc.encode_device(d, offs, PYTWIN_KITTI, want_front=False)
    assert f is None and t is not None
    t, f = enc.encode_device(d, offs, PYTWIN_KITTI, want_top=False)
    assert t is None and f is not None
    assert enc.launch_count > 0
    enc.close()


def test_device_atan2f_equals_libm(pytwin):
    """The device restatement of glibc's float atan2f, bit for bit, on 16 M arguments incl. special values."""
    import torch
    from mv3d_b200.pytwin import PyTwinEncoder
    enc = PyTwinEncoder(0)
    rng = np.random.default_rng(5)
    bad = total = 0
    for scale in (80.0, 2.0, 1e-3, 1e5):
        x = rng.uniform(-scale, scale, 4_000_000).astype(np.float32)
        y = rng.uniform(-scale, scale, 4_000_000).astype(np.float32)
        x[:12] = [0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, 0.0, 5.0, -5.0, 1.0, 1e-30, 3e30]
        y[:12] = [0.0, 0.0, 0.0, -0.0, 1.0, np.inf, -3.0, np.inf, -np.inf, 7.0, 2e30, 1e-30]
        got = enc.debug_atan2f(torch.from_numpy(y).cuda(), torch.from_numpy(x).cuda()).cpu().numpy()
        exp = pytwin.atan2f(y, x, "libm")
        bad += np.count_nonzero(got.view(np.uint32) != exp.view(np.uint32))
        total += len(x)
    assert bad == 0, f"{bad} of {total} atan2f results differ from the C library"
    enc.close()


def test_batch16_every_frame(pytwin):
    """16 KITTI-shaped sweeps in one batch, every frame against the oracle."""
    import torch
    from mv3d_b200.pytwin import PyTwinEncoder
    enc = PyTwinEncoder(0)
    sweeps = [hdl64_sweep(f) for f in range(40, 56)]
    pts, offs = pack_sweeps(sweeps)
    top, front = enc.encode_device(torch.from_numpy(pts).cuda(), offs, PYTWIN_KITTI)
    torch.cuda.synchronize()
    top, front = top.cpu().numpy(), front.cpu().numpy()
    for i, s in enumerate(sweeps):
        same(top[i], pytwin.lidar_to_top(s, PYTWIN_KITTI), f"frame {i} top")
        same(front[i], pytwin.lidar_to_front_fast(s, PYTWIN_KITTI), f"frame {i} front")
    enc.close()


@pytest.mark.parametrize("readback", ["compact", "dense"])
def test_host_batch_readback(pytwin, readback, monkeypatch):
    """numpy in / numpy out for a ragged batch, through both ways of bringing the bird's-eye view back: the compact
    list of touched cells rebuilt by host threads (default) and the dense device-to-host copy."""
    from mv3d_b200.pytwin import PyTwinEncoder
    monkeypatch.setenv("MV3D_HOST_READBACK", readback)
    enc = PyTwinEncoder(0)
    qc = {c[0]: c[1] for c in pytwin_cases.quirk_cases()}
    sweeps = [hdl64_sweep(6), np.zeros((0, 4), np.float32), qc["integral_qz"], hdl64_sweep(7)[:30001], hdl64_sweep(8)[:1],
              qc["ties_density"], hdl64_sweep(9)[::3]]
    pts, offs = pack_sweeps(sweeps)
    for p in (PYTWIN_KITTI, PYTWIN_DIDI2):
        top, front = enc.encode_host(pts, offs, p)
        for i, s in enumerate(sweeps):
            same(top[i], pytwin.lidar_to_top(s, p), f"{readback} sweep {i} top")
            same(front[i], pytwin.lidar_to_front_fast(s, p), f"{readback} sweep {i} front")
    top, front = enc.encode_host(pts, offs, PYTWIN_KITTI, want_front=False)
    assert front is None
    same(top[0], pytwin.lidar_to_top(sweeps[0], PYTWIN_KITTI), "top only")
    h2d, d2h = ctypes_transfer(enc)
    assert h2d > 0 and d2h > 0
    enc.close()


def ctypes_transfer(enc):
    import ctypes
    h2d, d2h = ctypes.c_ulonglong(), ctypes.c_ulonglong()
    assert enc._lib.mv3d_ctx_transfer_bytes(enc._ctx, ctypes.byref(h2d), ctypes.byref(d2h), 0) == 0
    return h2d.value, d2h.value


def test_loader_names_pass_the_reference_self_test(pytwin):
    """The reference's own equality asserts for its pycuda twins (src/net/utility/front_top_preprocess.py:211-223),
    run against the drop-in names the loaders import, with the executed-reference golden outputs as ground truth."""
    from mv3d_b200.pytwin import Preprocess, lidar_to_front_cuda, lidar_to_top_cuda
    g = np.load(GOLDEN)
    name, lidar, p = next(c for c in CASES if c[2] is PYTWIN_KITTI and len(c[1]) > 1000)
    pro = Preprocess()
    top = lidar_to_top_cuda(lidar)
    top_gt = unpack(g, name, "top", p.top_shape)
    assert top.shape == top_gt.shape
    zn = p.top_shape[2] - 2
    assert np.sum(top[..., 0:zn] != top_gt[..., 0:zn]) == 0
    assert np.sum(top[..., zn + 1] != top_gt[..., zn + 1]) == 0
    assert np.sum(top[..., zn] != top_gt[..., zn]) == 0  # the intensity too, which the pycuda twin does not promise
    front = lidar_to_front_cuda(lidar)
    front_gt = unpack(g, name, "front", p.front_shape)
    assert front.shape == front_gt.shape
    assert np.sum(front != front_gt) == 0
    same(pro.lidar_to_top(lidar), top_gt, "Preprocess.lidar_to_top")
    same(pro.lidar_to_front_fast(lidar), front_gt, "Preprocess.lidar_to_front_fast")
