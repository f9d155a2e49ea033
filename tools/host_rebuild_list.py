"""A read-back stream shaped like tile_compact_kernel's for one synthetic HDL-64E sweep (input of
tools/host_rebuild_bench.cu): u32 n_words, tiles, tile_cells; table[2*tiles+1] = {first word, cells} per tile + n_words;
words[n_words] = per touched cell {cell, plane mask, non-zero values} (a tile's cells ascending)."""
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
keep = (X >= 1) & (X <= 704) & (Y >= 1) & (Y <= 800) & (Z >= 0) & (Z < 25)
T, planes = 340, 27
tiles = (704 * 800 + T - 1) // T
cell = ((X - 1) * 800 + (Y - 1))[keep]
uc, inv = np.unique(cell, return_inverse=True)  # ascending: tile by tile, cells ascending inside a tile
mask = np.full(len(uc), (1 << 25) | (1 << 26), np.uint32)  # intensity + density planes
np.bitwise_or.at(mask, inv, (1 << Z[keep]).astype(np.uint32))
nnz = np.array([bin(int(m)).count("1") for m in mask])
per_tile = np.bincount(uc // T, minlength=tiles)
words_per_cell = 2 + nnz
first_word = np.concatenate([[0], np.cumsum(words_per_cell)[:-1]])
n_words = int(words_per_cell.sum())
words = np.full(n_words, 0x3F800000, np.uint32)
words[first_word] = uc
words[first_word + 1] = mask
first_cell_of_tile = np.concatenate([[0], np.cumsum(per_tile)[:-1]])
table = np.zeros(2 * tiles + 1, np.uint32)
table[0:2 * tiles:2] = np.where(per_tile > 0, first_word[np.minimum(first_cell_of_tile, len(uc) - 1)], 0)
table[1:2 * tiles:2] = per_tile
table[2 * tiles] = n_words
print('kept points', int(keep.sum()), 'touched cells', len(uc), 'non-zero planes per cell %.2f' % nnz.mean(),
      'stream', n_words * 4, 'bytes, tiles', tiles)
with open(sys.argv[1] if len(sys.argv) > 1 else '/tmp/list.bin', 'wb') as f:
    np.array([n_words, tiles, T], np.uint32).tofile(f)
    table.tofile(f)
    words.tofile(f)
