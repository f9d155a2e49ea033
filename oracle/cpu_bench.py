"""Times the CPU implementation of the view encoder on the host cores (TEST/BENCH INFRASTRUCTURE ONLY).

Used by bench.py's `cpu_baseline` leg and by `bench.py --impl reference`. Fan-out is the reference's own
idiom: one `multiprocessing.Pool` of worker processes, one sweep per task (`src/data.py:495-506`,
`src/preprocess_for_mv3d.py:62-72`). `kind` is "reference" when oracle/_ref (the reference's unmodified C
sources, built by oracle/Makefile) is present, else "port" (oracle/view_oracle.c).
"""
from __future__ import annotations

import multiprocessing as mp
import os
import time

_STATE = {}


def _init(kind, top, front, want_front):
    import oracle
    _STATE.update(kind=kind, top=top, front=front, want_front=want_front, oracle=oracle, cache={})


def _encode(frame_id):
    from mv3d_b200.sweeps import hdl64_sweep
    st = _STATE
    pts = st["cache"].get(frame_id)
    if pts is None:
        pts = st["cache"].setdefault(frame_id, hdl64_sweep(frame_id))
    impl = "reference" if st["kind"] == "reference" else "restated"
    t0 = time.perf_counter()
    if st["want_front"]:
        st["oracle"].top_and_front_maps(pts, st["top"], st["front"], impl)
    else:
        st["oracle"].top_maps(pts, st["top"], impl)
    return time.perf_counter() - t0


class CpuEncoderPool:
    """A pool of `cores` processes, each calling the CPU routine on whole sweeps."""

    def __init__(self, top, front, cores=None, want_front=True):
        import oracle
        oracle.build()
        self.kind = "reference" if oracle.reference_available() else "port"
        self.cores = int(cores or os.cpu_count() or 1)
        # One call in THIS process before the workers fork: loads the timed library (oracle/_ref/*.so for kind
        # "reference") into the parent too, so whoever watches which shared objects the benchmark process maps sees
        # it; the workers inherit the mapping.
        import numpy as np
        impl = "reference" if self.kind == "reference" else "restated"
        warm = np.array([[10.0, 1.0, -1.0, 0.5]], np.float32)
        if want_front:
            oracle.top_and_front_maps(warm, top, front, impl)
        else:
            oracle.top_maps(warm, top, impl)
        ctx = mp.get_context("fork")  # create BEFORE CUDA is initialised in the parent
        self.pool = ctx.Pool(self.cores, initializer=_init, initargs=(self.kind, top, front, want_front))

    def run(self, frame_ids):
        """Encodes the given sweeps across the pool; returns (wall seconds, per-call seconds)."""
        chunk = max(1, len(frame_ids) // (self.cores * 4))
        t0 = time.perf_counter()
        per_call = self.pool.map(_encode, list(frame_ids), chunksize=chunk)
        return time.perf_counter() - t0, per_call

    def close(self):
        self.pool.close()
        self.pool.join()
