"""Randomised parity (GPU): random grids, resolutions, front-view windows and clouds through every strategy of the
batched device entry point and through the two reference symbols, against the CPU oracle, bit for bit."""
import numpy as np
import pytest

import cases
from mv3d_b200.sweeps import pack_sweeps

pytestmark = pytest.mark.gpu
FUZZ = cases.fuzz_cases()


def _same(a, b):
    return a.shape == b.shape and np.array_equal(a, b, equal_nan=True)


@pytest.mark.parametrize("name,pts,top,front", FUZZ, ids=[c[0] for c in FUZZ])
def test_fuzz_device_all_strategies(oracle_mod, name, pts, top, front):
    import torch
    from mv3d_b200.encoder import ViewEncoder
    sweeps = [pts, pts[::-1].copy(), pts[: len(pts) // 3]]  # reversed order flips every tie-break
    packed, offs = pack_sweeps(sweeps)
    d = torch.from_numpy(packed).cuda() if len(packed) else torch.zeros((0, 4), device="cuda")
    for semantics in ("top", "topfront"):
        want = [oracle_mod.top_and_front_maps(s, top, front) if semantics == "topfront" else (oracle_mod.top_maps(s, top), None)
                for s in sweeps]
        for strategy in ("auto", "tile", "atomic", "sort", "patch"):
            enc = ViewEncoder(0, semantics=semantics, strategy=strategy)
            t, f = enc.encode_device(d, offs, top, front if semantics == "topfront" else None)
            torch.cuda.synchronize()
            for i in range(len(sweeps)):
                assert _same(t[i].cpu().numpy(), want[i][0]), f"{name} {semantics}/{strategy} sweep {i}: top differs"
                if semantics == "topfront":
                    assert _same(f[i].cpu().numpy(), want[i][1]), f"{name} {semantics}/{strategy} sweep {i}: front differs"
            enc.close()


@pytest.mark.parametrize("name,pts,top,front", FUZZ, ids=[c[0] for c in FUZZ])
def test_fuzz_reference_symbols(oracle_mod, name, pts, top, front):
    from mv3d_b200.ver3 import createTopAndFrontMaps, createTopMaps
    out = np.full(top.shape, 5.0, np.float32)
    createTopMaps(pts, pts.shape[0], out, top.as_top_paras())
    assert _same(out, oracle_mod.top_maps(pts, top)), f"{name}: createTopMaps differs"
    out = np.full(top.shape, 5.0, np.float32)
    fout = np.full(front.shape, 5.0, np.float32)
    createTopAndFrontMaps(pts, pts.shape[0], out, fout, top.as_top_paras(), front.as_front_paras())
    et, ef = oracle_mod.top_and_front_maps(pts, top, front)
    assert _same(out, et) and _same(fout, ef), f"{name}: createTopAndFrontMaps differs"
