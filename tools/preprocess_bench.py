#!/usr/bin/env python
"""End-to-end rate of the preprocessing run (mv3d_b200/preprocess.py = src/preprocess_for_mv3d.py on the GPU):
N KITTI-shaped `.bin` files -> top_view/<tag>.npy (800x600x27) + front_view/<tag>.npy (1500x100x3), files in tmpfs so
that the numbers are about the software, not the disk. Also times the reference's per-frame file I/O alone
(np.fromfile + two np.save calls, one thread) and the compressed-archive writer. One JSON line per case.
python tools/preprocess_bench.py [--frames 64] [--batch 16]"""
import argparse
import json
import os
import shutil
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from mv3d_b200 import sweep_io
from mv3d_b200.params import PYTWIN_KITTI as P
from mv3d_b200.preprocess import preprocess_directory
from mv3d_b200.pytwin import PyTwinEncoder
from mv3d_b200.sweeps import hdl64_sweep

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=64)
ap.add_argument("--batch", type=int, default=16)
ap.add_argument("--io-threads", type=int, default=0)
args = ap.parse_args()

base = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
root = tempfile.mkdtemp(prefix="mv3d_pre_", dir=base)
try:
    os.makedirs(os.path.join(root, "velodyne"))
    for f in range(args.frames):
        hdl64_sweep(f).tofile(os.path.join(root, "velodyne", f"{f:06d}.bin"))
    enc = PyTwinEncoder(0)
    preprocess_directory(root, P, batch=args.batch, io_threads=args.io_threads, encoder=enc)   # warm-up: contexts, pages
    stats = preprocess_directory(root, P, batch=args.batch, io_threads=args.io_threads, encoder=enc)
    out_bytes = 4 * (int(np.prod(P.top_shape)) + int(np.prod(P.front_shape)))
    print(json.dumps({"case": "preprocess_run", "frames": stats["frames"], "batch": args.batch,
                      "frames_per_s": round(stats["frames"] / stats["total_s"], 1),
                      "ms_per_frame": round(1e3 * stats["total_s"] / stats["frames"], 2),
                      "read_s": round(stats["read_s"], 3), "encode_s": round(stats["encode_s"], 3),
                      "write_s": round(stats["write_s"], 3), "total_s": round(stats["total_s"], 3),
                      "written_GBps": round(stats["frames"] * out_bytes / stats["write_s"] / 1e9, 2),
                      "tmpfs": base is not None}))

    # the reference's file I/O per frame, alone (no encoding): np.fromfile + np.save x 2, single thread
    top = sweep_io.load_view(os.path.join(root, "top_view", "000000.npy"))
    front = sweep_io.load_view(os.path.join(root, "front_view", "000000.npy"))
    n = min(8, args.frames)
    t0 = time.perf_counter()
    for f in range(n):
        scan = np.fromfile(os.path.join(root, "velodyne", f"{f:06d}.bin"), dtype=np.float32).reshape((-1, 4))
        np.save(os.path.join(root, "np_top.npy"), top)
        np.save(os.path.join(root, "np_front.npy"), front)
    np_s = (time.perf_counter() - t0) / n
    print(json.dumps({"case": "numpy_io_only", "ms_per_frame": round(1e3 * np_s, 2), "threads": 1}))

    # compressed archives: np.savez_compressed (src/data.py:521) vs the batched native writer
    views = np.stack([sweep_io.load_view(os.path.join(root, "top_view", f"{f:06d}.npy")) for f in range(min(16, args.frames))])
    t0 = time.perf_counter()
    np.savez_compressed(os.path.join(root, "np_z.npy"), top_view=views[0])
    z_np = time.perf_counter() - t0
    paths = [os.path.join(root, f"z{f:03d}.npy") for f in range(views.shape[0])]
    t0 = time.perf_counter()
    sweep_io.savez_compressed_batch(paths, views, threads=args.io_threads)
    z_mine = (time.perf_counter() - t0) / views.shape[0]
    print(json.dumps({"case": "savez_compressed", "numpy_ms_per_frame": round(1e3 * z_np, 1),
                      "batched_ms_per_frame": round(1e3 * z_mine, 1), "frames": int(views.shape[0]),
                      "archive_bytes": os.path.getsize(paths[0] + ".npz"),
                      "numpy_archive_bytes": os.path.getsize(os.path.join(root, "np_z.npy.npz"))}))
finally:
    shutil.rmtree(root, ignore_errors=True)
