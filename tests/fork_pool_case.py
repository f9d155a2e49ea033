"""Run as a script by tests/test_gpu_concurrency.py in a FRESH interpreter: the reference's fan-out idiom
(`multiprocessing.Pool(3).map` over frames, src/data.py:495-506; `Pool(12)` in src/preprocess_for_mv3d.py:62-72) with
fork workers whose first CUDA use happens after the fork. The parent never touches CUDA. Exits 0 when every frame
equals the oracle bit for bit."""
import multiprocessing
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np

from mv3d_b200.params import KITTI_FRONT_64x512 as FRONT, KITTI_PAPER_TOP as TOP
from mv3d_b200.sweeps import hdl64_sweep


def handler(frame):
    from mv3d_b200.ver3 import createTopAndFrontMaps, createTopMaps  # first CUDA use: in the worker
    s = hdl64_sweep(frame)
    top = np.ones(TOP.shape, np.float32)
    front = np.ones(FRONT.shape, np.float32)
    if frame % 2:
        createTopAndFrontMaps(s, s.shape[0], top, front, TOP.as_top_paras(), FRONT.as_front_paras())
    else:
        createTopMaps(s, s.shape[0], top, TOP.as_top_paras())
        front = None
    return frame, os.getpid(), top, front


def main():
    frames = list(range(40, 49))
    ctx = multiprocessing.get_context("fork")
    with ctx.Pool(3) as pool:
        results = pool.map(handler, frames, chunksize=1)
    import oracle  # the checker, in the parent only
    oracle.build()
    pids = set()
    for frame, pid, top, front in results:
        pids.add(pid)
        s = hdl64_sweep(frame)
        if frame % 2:
            et, ef = oracle.top_and_front_maps(s, TOP, FRONT)
            assert np.array_equal(top, et) and np.array_equal(front, ef), f"frame {frame} differs"
        else:
            assert np.array_equal(top, oracle.top_maps(s, TOP)), f"frame {frame} differs"
    assert len(pids) >= 2, "the pool did not fan out"
    print(f"fork pool ok: {len(frames)} frames on {len(pids)} workers")


if __name__ == "__main__":
    main()
