"""CPU tests of the host half of the compact read-back (mv3d_debug_expand_compact): the dense tensor rebuilt from
{cell, values} entries must equal a numpy scatter, for every thread count, alignment and edge case."""
import numpy as np
import pytest


def _expand(lib, cells, planes, idx, vals, threads, offset_floats=0, garbage=7.0, split=False):
    buf = np.full(cells * planes + offset_floats + 8, garbage, dtype=np.float32)
    dst = buf[offset_floats:offset_floats + cells * planes]
    entries = np.empty((len(idx), planes + 1), dtype=np.uint32)
    entries[:, 0] = idx
    entries[:, 1:] = vals.view(np.uint32)
    fn = lib.mv3d_debug_expand_split if split else lib.mv3d_debug_expand_compact
    rc = fn(dst.ctypes.data, cells, planes, entries.ctypes.data if len(idx) else None, len(idx), threads)
    assert rc == 0
    # nothing outside the tensor is touched
    assert (buf[:offset_floats] == garbage).all() and (buf[offset_floats + cells * planes:] == garbage).all()
    return dst.reshape(cells, planes)


@pytest.mark.parametrize("threads", [1, 2, 3, 8])
@pytest.mark.parametrize("cells,planes", [(563200, 27), (1000, 10), (7, 3), (1, 27), (300, 700), (5000, 1)])
def test_expand_matches_numpy_scatter(encoder_lib, threads, cells, planes):
    rng = np.random.default_rng(cells + threads)
    n = min(cells, 25000)
    idx = rng.choice(cells, size=n, replace=False).astype(np.uint32)
    vals = rng.random((n, planes), dtype=np.float32) + 0.5
    expect = np.zeros((cells, planes), np.float32)
    expect[idx] = vals
    for offset in (0, 1, 3):  # 16-byte aligned and misaligned destinations
        got = _expand(encoder_lib, cells, planes, idx, vals, threads, offset_floats=offset)
        assert np.array_equal(got, expect)
        # the two-step form the first sweep of a host call takes (zero-fill during the GPU's work, then scatter)
        got = _expand(encoder_lib, cells, planes, idx, vals, threads, offset_floats=offset, split=True)
        assert np.array_equal(got, expect)


def test_expand_empty_and_out_of_range(encoder_lib):
    got = _expand(encoder_lib, 100, 5, np.zeros(0, np.uint32), np.zeros((0, 5), np.float32), 4)
    assert not got.any()
    idx = np.array([3, 100, 0xFFFFFFFF, 99], np.uint32)  # 100 and 0xFFFFFFFF are out of range: ignored
    vals = np.arange(20, dtype=np.float32).reshape(4, 5) + 1
    expect = np.zeros((100, 5), np.float32)
    expect[3], expect[99] = vals[0], vals[3]
    for split in (False, True):
        assert np.array_equal(_expand(encoder_lib, 100, 5, idx, vals, 4, split=split), expect)
        assert not _expand(encoder_lib, 100, 5, np.zeros(0, np.uint32), np.zeros((0, 5), np.float32), 4, split=split).any()


def test_expand_preserves_nan_and_bit_patterns(encoder_lib):
    vals = np.array([[np.nan, -0.0, np.inf, 1e-45]], np.float32)
    got = _expand(encoder_lib, 10, 4, np.array([5], np.uint32), vals, 2)
    assert np.array_equal(got[5].view(np.uint32), vals[0].view(np.uint32))
