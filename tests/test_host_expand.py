"""CPU tests of the host half of the compact read-back (mv3d_debug_expand_sorted / _split): the dense tensor rebuilt from
the stream the compact tile kernel reports -- per touched cell {cell index, plane masks, non-zero values}, a table
{first word, touched cells} per tile -- must equal a numpy scatter, for every thread count, alignment and edge case,
and nothing may ever be written outside the destination, whatever the stream holds."""
import numpy as np
import pytest


def _make_stream(rng, cells, planes, tile_cells, fill, density, spare=3):
    """A stream laid out as tile_compact_kernel writes it: every tile's touched cells ascending in a range of its own,
    the ranges in the order the tiles happened to finish (shuffled here) with junk words between them."""
    tiles = (cells + tile_cells - 1) // tile_cells
    groups = (planes + 31) // 32
    touched = np.sort(rng.choice(cells, size=int(cells * fill), replace=False)) if fill > 0 else np.zeros(0, np.int64)
    expect = np.zeros((cells, planes), np.float32)
    nonzero = rng.random((len(touched), planes)) < density
    nonzero[:, planes - 1] = True  # a touched cell always has a non-zero density plane
    values = (rng.random((len(touched), planes), dtype=np.float32) + 0.5) * nonzero
    expect[touched] = values
    per_tile = [[] for _ in range(tiles)]
    for j, c in enumerate(touched):
        masks = [0] * groups
        for k in np.flatnonzero(nonzero[j]):
            masks[k // 32] |= 1 << (k % 32)
        per_tile[c // tile_cells].append(np.concatenate([np.array([c] + masks, np.uint32),
                                                         values[j][nonzero[j]].view(np.uint32)]))
    words, table = [], np.zeros(2 * tiles + 1, np.uint32)
    pos = 0
    for t in rng.permutation(tiles):
        gap = int(rng.integers(0, spare + 1))
        words.append(rng.integers(0, 2 ** 32, gap, dtype=np.uint64).astype(np.uint32))  # junk the table never points at
        pos += gap
        table[2 * t], table[2 * t + 1] = pos, len(per_tile[t])
        for e in per_tile[t]:
            words.append(e)
            pos += len(e)
    words = np.concatenate(words) if words else np.zeros(0, np.uint32)
    table[2 * tiles] = len(words)
    return np.ascontiguousarray(words), table, tiles, expect


def _run(lib, fn, cells, planes, words, table, tiles, tile_cells, threads, offset=0, garbage=7.0):
    buf = np.full(cells * planes + offset + 40, garbage, dtype=np.float32)
    dst = buf[offset:offset + cells * planes]
    keep = words if len(words) else np.zeros(1, np.uint32)
    rc = getattr(lib, fn)(dst.ctypes.data, cells, planes, keep.ctypes.data, len(words), table.ctypes.data, tiles, tile_cells, threads)
    assert rc == 0
    assert (buf[:offset] == garbage).all() and (buf[offset + cells * planes:] == garbage).all(), "wrote outside the tensor"
    return dst.reshape(cells, planes)


@pytest.mark.parametrize("threads", [1, 3, 8])
@pytest.mark.parametrize("cells,planes,tile_cells,fill,density",
                         [(563200, 27, 340, 0.05, 0.2), (563200, 27, 352, 0.6, 0.15), (5000, 10, 64, 1.0, 1.0),
                          (1000, 1, 7, 0.3, 1.0), (999, 3, 1000, 0.5, 0.5), (77, 27, 340, 0.2, 0.3), (4096, 16, 32, 0.0, 0.5),
                          (3000, 40, 50, 0.3, 0.3), (300, 700, 10, 0.3, 0.05)])
def test_rebuild_matches_numpy_scatter(encoder_lib, threads, cells, planes, tile_cells, fill, density):
    """Sparse, dense, every cell touched with every plane set, one-plane cells, a single tile, more than 32 planes (two
    mask words), cells wider than the compose window (zero-then-place path); aligned and misaligned destinations; the
    single pass and the two-step form."""
    rng = np.random.default_rng(cells * 31 + threads)
    words, table, tiles, expect = _make_stream(rng, cells, planes, tile_cells, fill, density)
    for offset in (0, 1, 5, 16):
        for fn in ("mv3d_debug_expand_sorted", "mv3d_debug_expand_split"):
            got = _run(encoder_lib, fn, cells, planes, words, table, tiles, tile_cells, threads, offset)
            assert np.array_equal(got.view(np.uint32), expect.view(np.uint32)), (fn, offset)


def test_rebuild_empty_stream(encoder_lib):
    table = np.zeros(2 * 7 + 1, np.uint32)
    for fn in ("mv3d_debug_expand_sorted", "mv3d_debug_expand_split"):
        assert not _run(encoder_lib, fn, 100, 5, np.zeros(0, np.uint32), table, 7, 16, 4).any()


def test_rebuild_preserves_bit_patterns(encoder_lib):
    """NaN, -0, inf and denormals travel as bits; -0 counts as a non-zero plane (its bits are not zero)."""
    vals = np.array([np.nan, -0.0, np.inf, 1e-45], np.float32)
    words = np.concatenate([np.array([5, 0b1111], np.uint32), vals.view(np.uint32)])
    table = np.array([0, 1, len(words)], np.uint32)
    for fn in ("mv3d_debug_expand_sorted", "mv3d_debug_expand_split"):
        got = _run(encoder_lib, fn, 10, 4, words, table, 1, 10, 2)
        assert np.array_equal(got[5].view(np.uint32), vals.view(np.uint32))
        assert not got[:5].any() and not got[6:].any()


def test_rebuild_never_trusts_the_stream(encoder_lib):
    """Cells out of order, repeated, of another tile or out of range are skipped; mask bits beyond the planes are ignored
    (their values are still consumed); an entry or a table range running past the stream ends its tile; a table pointing
    anywhere is harmless -- and in every case nothing is written outside the destination."""
    cells, planes, tile_cells = 64, 5, 16
    F = lambda *v: list(np.array(v, np.float32).view(np.uint32))  # noqa: E731
    e = lambda cell, mask, *v: [cell, mask] + F(*v)  # noqa: E731
    t0 = e(3, 0b00001, 1.0) + e(2, 0b00001, 2.0) + e(3, 0b00010, 3.0) + e(15, 0b10001, 4.0, 5.0)   # 2 and the second 3: skipped
    t1 = e(40, 0b1, 6.0) + e(16, 0b1100000, 7.0, 8.0) + e(17, 0b100011, 9.0, 10.0, 11.0)            # 40: another tile's; bits 5,6: ignored
    t3 = e(48, 0b1, 12.0) + e(63, 0b11111, 13.0, 14.0, 15.0, 16.0, 17.0) + e(64, 0b1, 18.0) + e(5, 0b1, 19.0)  # 64: no cell
    t2 = e(33, 0b1, 20.0) + [34, 0b11111] + F(21.0, 22.0)                                           # last entry runs past the stream
    words = np.array(t0 + t1 + t3 + t2, np.uint32)
    table = np.array([0, 4, len(t0), 3, len(t0 + t1 + t3), 100, len(t0 + t1), 4, len(words)], np.uint32)
    expect = np.zeros((cells, planes), np.float32)
    expect[3, 0], expect[15, 0], expect[15, 4] = 1.0, 4.0, 5.0
    expect[17, 0], expect[17, 1] = 9.0, 10.0          # cell 16 keeps no value: both its bits are beyond the planes
    expect[48, 0], expect[63] = 12.0, [13.0, 14.0, 15.0, 16.0, 17.0]
    expect[33, 0] = 20.0
    for threads in (1, 4):
        for fn in ("mv3d_debug_expand_sorted", "mv3d_debug_expand_split"):
            got = _run(encoder_lib, fn, cells, planes, words, table, 4, tile_cells, threads, offset=8, garbage=9.0)
            assert np.array_equal(got, expect), fn
    wild = np.array([10 ** 9, 5, 0xFFFFFFFF, 0xFFFFFFFF, 3, 2 ** 31, 0, 0, len(words)], np.uint32)
    for fn in ("mv3d_debug_expand_sorted", "mv3d_debug_expand_split"):
        _run(encoder_lib, fn, cells, planes, words, wild, 4, tile_cells, 3, offset=8)
    rng = np.random.default_rng(5)
    for _ in range(20):  # pure noise as stream and table
        noise = rng.integers(0, 2 ** 32, 4000, dtype=np.uint64).astype(np.uint32)
        noise[::7] = rng.integers(0, cells, len(noise[::7]))
        tab = rng.integers(0, 5000, 9, dtype=np.uint64).astype(np.uint32)
        for fn in ("mv3d_debug_expand_sorted", "mv3d_debug_expand_split"):
            _run(encoder_lib, fn, cells, planes, noise, tab, 4, tile_cells, 3, offset=3)
