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


def _sorted_list(rng, cells, planes, tile_cells, fill, spare):
    """A list laid out as tile_compact_kernel writes it: every tile's touched cells ascending in a range of its own, the
    ranges in the order the tiles happened to finish (shuffled here), a table {first entry, entries} per tile."""
    tiles = (cells + tile_cells - 1) // tile_cells
    touched = np.sort(rng.choice(cells, size=int(cells * fill), replace=False)) if fill > 0 else np.zeros(0, np.int64)
    per_tile = np.bincount(touched // tile_cells, minlength=tiles)
    order = rng.permutation(tiles)
    gaps = rng.integers(0, spare + 1, tiles)  # junk between ranges: never referenced by the table, must land nowhere
    first = np.zeros(tiles, np.int64)
    pos = 0
    for t in order:
        pos += int(gaps[t])
        first[t] = pos
        pos += int(per_tile[t])
    n = pos + 2
    table = np.zeros(2 * tiles + 1, np.uint32)
    table[0:2 * tiles:2] = first
    table[1:2 * tiles:2] = per_tile
    table[2 * tiles] = n
    idx = rng.integers(0, cells, n).astype(np.uint32)  # junk everywhere the table does not point
    vals = rng.random((n, planes), dtype=np.float32) + 0.5
    start_of = np.searchsorted(touched // tile_cells, np.arange(tiles))
    where = first[touched // tile_cells] + (np.arange(len(touched)) - start_of[touched // tile_cells])
    idx[where] = touched
    expect = np.zeros((cells, planes), np.float32)
    expect[touched] = vals[where]
    return idx, vals, table, tiles, expect


@pytest.mark.parametrize("threads", [1, 3, 8])
@pytest.mark.parametrize("cells,planes,tile_cells,fill", [(563200, 27, 340, 0.05), (563200, 27, 352, 0.6), (5000, 10, 64, 1.0),
                                                          (1000, 1, 7, 0.3), (999, 3, 1000, 0.5), (77, 27, 340, 0.2),
                                                          (4096, 16, 32, 0.0), (300, 700, 10, 0.3)])
def test_expand_sorted_matches_numpy_scatter(encoder_lib, threads, cells, planes, tile_cells, fill):
    """The linear-walk rebuild (what the host entry points run) on lists shaped like the compact tile kernel's:
    sparse, dense, every cell touched, one-plane cells, a single tile, cells wider than the compose window (falls back),
    aligned and misaligned destinations; guard floats on both sides stay untouched."""
    rng = np.random.default_rng(cells * 31 + threads)
    idx, vals, table, tiles, expect = _sorted_list(rng, cells, planes, tile_cells, fill, spare=3)
    if planes > 200:  # the fallback scatters every entry of the list: give it a list without junk
        keep = np.concatenate([np.arange(table[2 * t], table[2 * t] + table[2 * t + 1]) for t in range(tiles)]).astype(np.int64)
        idx, vals = idx[keep].copy(), vals[keep].copy()
        table[0:2 * tiles:2] = np.concatenate([[0], np.cumsum(table[1:2 * tiles:2])[:-1]])
    for offset in (0, 1, 5, 16):
        buf = np.full(cells * planes + offset + 40, 7.0, dtype=np.float32)
        dst = buf[offset:offset + cells * planes]
        rc = encoder_lib.mv3d_debug_expand_sorted(dst.ctypes.data, cells, planes, idx.ctypes.data, vals.ctypes.data, len(idx),
                                                  table.ctypes.data, tiles, tile_cells, threads)
        assert rc == 0
        assert (buf[:offset] == 7.0).all() and (buf[offset + cells * planes:] == 7.0).all()
        assert np.array_equal(dst.reshape(cells, planes).view(np.uint32), expect.view(np.uint32))


def test_expand_sorted_never_writes_out_of_contract_entries(encoder_lib):
    """An entry that breaks the ascending order, repeats a cell or names a cell of another tile is skipped, a range that
    runs past the list is clamped -- the walk never writes outside its own tile, whatever the list and table hold."""
    cells, planes, tile_cells = 64, 5, 16
    #              tile 0: slots 0-3      tile 1: slots 4-7           tile 3: slots 8-11      junk
    idx = np.array([3, 2, 3, 15,          40, 16, 17, 0xFFFFFFFF,     48, 63, 64, 5,          1, 2, 33, 34], np.uint32)
    table = np.array([0, 4,   4, 4,   14, 100,   8, 4,   16], np.uint32)  # tile 2: entries 14.. "100 of them" (list has 16)
    vals = (np.arange(16 * planes, dtype=np.float32).reshape(16, planes) + 1)
    expect = np.zeros((cells, planes), np.float32)
    for slot in (0, 3, 5, 6, 8, 9, 14, 15):  # 2 and the second 3 are out of order; 40 is another tile's; 64 is no cell
        expect[idx[slot]] = vals[slot]
    for threads in (1, 4):
        buf = np.full(cells * planes + 16, 9.0, np.float32)
        dst = buf[8:8 + cells * planes]
        assert encoder_lib.mv3d_debug_expand_sorted(dst.ctypes.data, cells, planes, idx.ctypes.data, vals.ctypes.data, 16,
                                                    table.ctypes.data, 4, tile_cells, threads) == 0
        assert (buf[:8] == 9.0).all() and (buf[8 + cells * planes:] == 9.0).all()
        assert np.array_equal(dst.reshape(cells, planes), expect)
