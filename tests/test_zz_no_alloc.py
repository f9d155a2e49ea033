"""GPU test of mv3d_front_view_noalloc (include/mv3d_encode.h): the allocation-free front view
(no counterpart in the reference). Named test_zz_* so that pytest collects it last:
a failure here must not hide the parity tests.

The new entry point is called through the C ABI (ctypes) and compared with the EXISTING entry point
(mv3d_encode_batch_dev with top_dev = NULL) on the same inputs, bit for bit -- the tolerance every front-view test of
this repository uses -- on two calls in a row that share the static scratch (different sweeps, so stale winners of the
first call would show in the second). It also checks the stated limit is enforced with a clear error."""
import ctypes

import numpy as np
import pytest

from mv3d_b200.params import KITTI_FRONT_64x512, KITTI_PAPER_TOP, FrontParams
from mv3d_b200.sweeps import hdl64_sweep, pack_sweeps

pytestmark = pytest.mark.gpu
MAX_PIXELS = 262144  # MV3D_NOALLOC_MAX_PIXELS


def _noalloc(lib, d_pts, offs, front_out, top_p, front_p, semantics):
    from mv3d_b200.encoder import _front_c, _top_c
    import torch
    o = np.ascontiguousarray(offs, dtype=np.int32)
    return lib.mv3d_front_view_noalloc(
        d_pts.data_ptr(), o.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), len(o) - 1, front_out.data_ptr(),
        ctypes.byref(_top_c(top_p)), ctypes.byref(_front_c(front_p)), semantics,
        ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))


@pytest.mark.parametrize("semantics", ["topfront", "top"])
def test_noalloc_front_view_equals_existing_entry_point(semantics):
    import torch
    assert torch.cuda.is_available(), "needs a CUDA device"
    from mv3d_b200 import _lib
    from mv3d_b200.encoder import ViewEncoder
    lib = _lib.load_library()
    sem = {"top": _lib.SEM_TOP, "topfront": _lib.SEM_TOPFRONT}[semantics]
    top_p, front_p = KITTI_PAPER_TOP, KITTI_FRONT_64x512
    enc = ViewEncoder(0, semantics=semantics)
    calls = [[hdl64_sweep(500 + i) for i in range(8)],                                  # 8 x 64 x 512 = the whole scratch
             [hdl64_sweep(520)[:70000], np.zeros((0, 4), np.float32), hdl64_sweep(521)]]  # ragged, with an empty sweep
    for sweeps in calls:
        pts, offs = pack_sweeps(sweeps)
        d = torch.from_numpy(pts).cuda()
        _, want = enc.encode_device(d, offs, top_p, front_p, want_top=False)
        got = torch.full_like(want, -5.0)  # every float of the output must be written
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        rc = _noalloc(lib, d, offs, got, top_p, front_p, sem)
        assert rc == 0, _lib.last_error()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        assert torch.equal(got, want), f"{semantics}: {int((got != want).sum())} floats differ from mv3d_encode_batch_dev"
        assert int((want != 0).sum()) > 1000 * (len(sweeps) - 1)  # not vacuous
    # a window of the packed buffer (offsets[0] > 0), same stream, right after: still equal
    pts, offs = pack_sweeps(calls[0])
    d = torch.from_numpy(pts).cuda()
    _, want = enc.encode_device(d, offs[2:6], top_p, front_p, want_top=False)
    got = torch.empty_like(want)
    assert _noalloc(lib, d, offs[2:6], got, top_p, front_p, sem) == 0, _lib.last_error()
    torch.cuda.synchronize()
    assert torch.equal(got, want)
    enc.close()


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    from mv3d_b200 import _lib
    lib = _lib.load_library()
    sweeps = [hdl64_sweep(0)[:1000]] * 9
    pts, offs = pack_sweeps(sweeps)
    d = torch.from_numpy(pts).cuda()
    out = torch.zeros((9, *KITTI_FRONT_64x512.shape), device="cuda")
    assert 9 * 64 * 512 > MAX_PIXELS
    assert _noalloc(lib, d, offs, out, KITTI_PAPER_TOP, KITTI_FRONT_64x512, _lib.SEM_TOPFRONT) != 0
    assert "MV3D_NOALLOC_MAX_PIXELS" in _lib.last_error() and "262144" in _lib.last_error()
    big = FrontParams(-20.0, 20.0, -45.0, 45.0, 0.05, 0.1, 800, 900, 3)  # one image of 720 000 pixels
    out = torch.zeros((1, *big.shape), device="cuda")
    assert _noalloc(lib, d, offs[:2], out, KITTI_PAPER_TOP, big, _lib.SEM_TOPFRONT) != 0
    assert "MV3D_NOALLOC_MAX_PIXELS" in _lib.last_error()
    assert not out.any().item()  # a refused call writes nothing
    assert _noalloc(lib, d, offs[:2], out[:, :64, :512].contiguous(), KITTI_PAPER_TOP, KITTI_FRONT_64x512, _lib.SEM_TOPFRONT) == 0
    torch.cuda.synchronize()
