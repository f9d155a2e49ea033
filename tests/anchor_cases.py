"""Seeded inputs for the empty-anchor filter tests (shared by the CPU and GPU suites)."""
import numpy as np


def sparse_view(shape, seed=0, fill=0.05, negative=False, integer=False):
    """A bird's-eye-view-like tensor: mostly zeros, `fill` of the cells hold values in every channel."""
    rng = np.random.default_rng(seed)
    xn, yn, c = shape
    view = np.zeros(shape, dtype=np.float32)
    hit = rng.random((xn, yn)) < fill
    k = int(hit.sum())
    if integer:
        vals = rng.integers(-3 if negative else 0, 4, size=(k, c)).astype(np.float32)
    else:
        vals = rng.random((k, c), dtype=np.float32)
        if negative:
            vals -= np.float32(0.5)
    view[hit] = vals
    return view


def dense_view(shape, seed=0):
    """`0.01*np.random.randn(W, H, C)` of the reference's own smoke test (remove_empty_box.py:64)."""
    rng = np.random.default_rng(seed)
    return (0.01 * rng.standard_normal(shape)).astype(np.float32)


def random_anchors(n, shape, seed=0, max_w=70, max_l=60, margin=30):
    """(y1, x1, y2, x2) boxes: most inside the view, some hanging over every edge, some inverted or empty,
    some higher than the 50 rows the reference's launch grid covers."""
    rng = np.random.default_rng(seed)
    xn, yn = shape[0], shape[1]
    x1 = rng.integers(-margin, xn + margin, size=n)
    y1 = rng.integers(-margin, yn + margin, size=n)
    w = rng.integers(-3, max_w, size=n)
    l = rng.integers(-3, max_l, size=n)
    return np.stack([y1, x1, y1 + l, x1 + w], axis=1).astype(np.int32)


def line_anchors(n, shape, seed=0, max_l=80):
    """Boxes exactly one row high (x2 = x1 + 1) inside the view: the reference kernel has no race on these."""
    rng = np.random.default_rng(seed)
    xn, yn = shape[0], shape[1]
    x1 = rng.integers(0, xn - 2, size=n)
    y1 = rng.integers(-5, yn - 1, size=n)
    l = rng.integers(0, max_l, size=n)
    return np.stack([y1, x1, y1 + l, x1 + 1], axis=1).astype(np.int32)


def smoke_anchors(n, seed=0):
    """The reference's smoke-test boxes (remove_empty_box.py:53-62): (0, 0, 40, 16) shifted by a random 0..900."""
    rng = np.random.default_rng(seed)
    shift = rng.integers(0, 901, size=(n, 1))
    return (np.array([[0, 0, 40, 16]]) + shift).astype(np.int32)
