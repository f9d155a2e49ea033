"""A touched-cell list shaped like tile_compact_kernel's for one synthetic HDL-64E sweep (input of tools/host_rebuild_bench.cu):
u32 n, tiles, tile_cells; table[2*tiles+1] = {first entry, entries} per tile + n; idx[n] (a tile's cells ascending)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mv3d_b200.sweeps import hdl64_sweep  # noqa: E402

s = hdl64_sweep(0)
x, y, z = s[:, 0], s[:, 1], s[:, 2]
X = np.floor((x - 0.0) / 0.1).astype(np.int64)
Y = np.floor((y + 40.0) / 0.1).astype(np.int64)
Z = np.floor((z + 4.2) / 0.2).astype(np.int64)
keep = (X >= 1) & (X <= 704) & (Y >= 1) & (Y <= 800) & (Z >= 0) & (Z <= 25)
T = 340
tiles = (704 * 800 + T - 1) // T
uc = np.unique(((X - 1) * 800 + (Y - 1))[keep])  # ascending: tile by tile, cells ascending inside a tile
per_tile = np.bincount(uc // T, minlength=tiles)
table = np.zeros(2 * tiles + 1, np.uint32)
table[1:2 * tiles:2] = per_tile
table[0:2 * tiles:2] = np.concatenate([[0], np.cumsum(per_tile)[:-1]])
table[2 * tiles] = len(uc)
print('kept points', int(keep.sum()), 'touched cells', len(uc), 'tiles', tiles)
with open(sys.argv[1] if len(sys.argv) > 1 else '/tmp/list.bin', 'wb') as f:
    np.array([len(uc), tiles, T], np.uint32).tofile(f)
    table.tofile(f)
    uc.astype(np.uint32).tofile(f)
