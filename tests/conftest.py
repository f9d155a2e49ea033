import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle_mod():
    """The CPU oracle (test infrastructure). Builds the restatement, and oracle/_ref where the reference exists."""
    import oracle
    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def encoder_lib():
    from mv3d_b200 import _lib
    return _lib.load_library()


@pytest.fixture(scope="module")
def torch_cuda():
    import torch
    assert torch.cuda.is_available(), "these tests need a CUDA device"
    return torch


_SELFCHECK = {"ctx": None, "enabled": None}


@pytest.fixture(autouse=True)
def _bounds_selfcheck(request):
    """On a library built with -DMV3D_SELF_CHECK (csrc/selfcheck.cuh: every index the kernels compute is tested; the
    stand-in for compute-sanitizer, which this GPU pool does not offer) no GPU test may leave a violation bit behind."""
    yield
    if request.node.get_closest_marker("gpu") is None or _SELFCHECK["enabled"] is False:
        return
    import ctypes
    from mv3d_b200 import _lib
    lib = _lib.load_library()
    flags = ctypes.c_uint(0)
    if _SELFCHECK["enabled"] is None:
        _SELFCHECK["enabled"] = lib.mv3d_debug_selfcheck(None, ctypes.byref(flags)) != 0  # 0: not compiled in
        if not _SELFCHECK["enabled"]:
            return
        ctx = ctypes.c_void_p()
        assert lib.mv3d_ctx_create(0, ctypes.byref(ctx)) == 0, _lib.last_error()
        _SELFCHECK["ctx"] = ctx
    assert lib.mv3d_debug_selfcheck(_SELFCHECK["ctx"], ctypes.byref(flags)) == 1, _lib.last_error()
    assert flags.value == 0, f"{request.node.nodeid}: kernel bounds-check violations 0x{flags.value:08x} (csrc/selfcheck.cuh)"
