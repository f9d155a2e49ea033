"""CPU tests of the boundary: libmv3d_encode.so builds for sm_100a, loads, and exports every symbol that
include/*.h declare -- and nothing on the product path reaches for the oracle."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = "".join(open(os.path.join(ROOT, "include", h)).read() for h in sorted(os.listdir(os.path.join(ROOT, "include"))))
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(createTopMaps|createTopAndFrontMaps|mv3d_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(encoder_lib):
    from mv3d_b200 import _lib
    declared = _declared_symbols()
    assert set(declared) == set(_lib.EXPORTED_SYMBOLS)
    for name in declared:
        assert hasattr(encoder_lib, name), name
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    for name in declared:
        assert re.search(rf" T {name}$", out, flags=re.M), f"{name} not exported as a C symbol"


def test_library_is_sm100a_only():
    from mv3d_b200 import _lib
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, out


def test_fails_loudly_without_gpu(encoder_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    ctx = ctypes.c_void_p()
    assert encoder_lib.mv3d_ctx_create(0, ctypes.byref(ctx)) != 0
    assert b"no CPU fallback" in encoder_lib.mv3d_last_error()
    import numpy as np
    from mv3d_b200.ver3 import createTopMaps
    from mv3d_b200.params import TopParams
    top = TopParams(0, 1, 0, 1, 0, 1, 0.5, 0.5, 0.5)
    out = np.ones(top.shape, np.float32)
    with pytest.raises(RuntimeError):
        createTopMaps(np.zeros((1, 4), np.float32), 1, out, top.as_top_paras())
    assert not out.any()  # failure leaves zeros, never stale data


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "mv3d_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(import|from)\s+oracle\b", text, flags=re.M), f
                assert "view_oracle" not in text and "oracle/_ref" not in text, f


def test_param_grid_sizes():
    from mv3d_b200.params import DENSE_TOP, KITTI_FRONT_64x512, KITTI_PAPER_TOP, REFERENCE_KITTI_TOP
    assert KITTI_PAPER_TOP.shape == (704, 800, 27)
    assert DENSE_TOP.shape == (1408, 1600, 10)
    assert REFERENCE_KITTI_TOP.shape == (800, 600, 27)  # src/utils/batch_loading.py:620
    assert KITTI_FRONT_64x512.shape == (64, 512, 3)
    assert KITTI_FRONT_64x512.as_front_paras()[6:] == (64, 512, 3)


REF_VER3 = "/root/reference/src/lidar_data_preprocess/Python_to_C_Interface/ver3"


@pytest.mark.skipif(not os.path.isdir(REF_VER3), reason="the reference tree is not on this machine")
def test_reference_wrappers_bind_this_library(encoder_lib, capfd):
    """The reference's OWN ctypes wrappers (ver3/createTopMaps.py, ver3/createTopAndFrontMaps.py), imported unmodified
    from where they lie and pointed at libmv3d_encode.so through their `c_lib_path` argument: the symbols resolve with
    the reference's call shape (15 / 25 by-value arguments, no argtypes), nothing crashes, and on a machine without a
    GPU the error contract holds (outputs zero-filled, the reason recorded)."""
    import importlib.util
    import numpy as np
    import torch
    from mv3d_b200 import _lib
    from mv3d_b200.params import FrontParams, TopParams

    def load(name):
        spec = importlib.util.spec_from_file_location("ref_" + name, os.path.join(REF_VER3, name + ".py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return getattr(mod, name)

    top = TopParams(0.0, 4.0, -2.0, 2.0, -1.0, 1.0, 0.5, 0.5, 0.5)
    front = FrontParams(-20.0, 20.0, -45.0, 45.0, 2.5, 5.0)
    pts = np.array([[1.1, 0.3, 0.2, 0.5], [2.2, -0.7, -0.4, 0.9]], np.float32)
    out = np.ones(top.shape, np.float32)
    load("createTopMaps")(pts, 2, out, list(top.as_top_paras()), _lib.LIB_PATH)
    fout = np.ones(front.shape, np.float32)
    out2 = np.ones(top.shape, np.float32)
    load("createTopAndFrontMaps")(pts, 2, out2, fout, list(top.as_top_paras()), list(front.as_front_paras()), _lib.LIB_PATH)
    if not torch.cuda.is_available():
        assert not out.any() and not out2.any() and not fout.any()
        assert b"no CPU fallback" in encoder_lib.mv3d_last_error()
    else:
        assert out.any() and out2.any()  # a GPU is present: the views were encoded


def test_strict_mode_aborts_instead_of_returning_zeros():
    """MV3D_STRICT=1: a failing reference symbol (here: no GPU / bad sizes) must not hand zeros to a caller that cannot
    tell them from a result -- it aborts (round-1 advisor finding). Without the switch it zero-fills and reports."""
    import sys
    code = (
        "import ctypes, numpy as np, sys\n"
        f"sys.path.insert(0, {ROOT!r})\n"
        "from mv3d_b200 import _lib\n"
        "lib = _lib.load_library()\n"
        "F, I = ctypes.c_float, ctypes.c_int\n"
        "pts = np.zeros((4, 4), np.float32); top = np.ones((2, 2, 3), np.float32); front = np.ones((4, 4, 3), np.float32)\n"
        "lib.createTopAndFrontMaps(ctypes.c_void_p(pts.ctypes.data), I(4), ctypes.c_void_p(top.ctypes.data),\n"
        "    ctypes.c_void_p(front.ctypes.data), F(0), F(2), F(0), F(2), F(0), F(1), F(1), F(1), F(1), I(2), I(2), I(1),\n"
        "    F(-1), F(1), F(-1), F(1), F(0.5), F(0.5), F(4.5), F(4), F(3))\n"  # C_SIZE = 4.5: refused with or without a GPU
        "print('returned', float(top.sum()), float(front.sum()))\n")
    env = dict(os.environ, MV3D_STRICT="1")
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env)
    assert res.returncode != 0 and "MV3D_STRICT is set, aborting" in res.stderr and "returned" not in res.stdout
    env.pop("MV3D_STRICT")
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env)
    assert res.returncode == 0 and "returned 0.0 0.0" in res.stdout and "createTopAndFrontMaps failed" in res.stderr


def test_host_fill_probe(encoder_lib):
    """The diagnostic bench.py prints beside the end-to-end figure runs without a GPU and gives a plausible rate."""
    gbs = encoder_lib.mv3d_debug_host_fill_gbs(64 << 20, 2, 2, 0, 0)
    assert 0.5 < gbs < 5000.0
    assert encoder_lib.mv3d_debug_host_fill_gbs(16, 2, 2, 0, 0) < 0
