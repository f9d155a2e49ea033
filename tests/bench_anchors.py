#!/usr/bin/env python
"""Measures the empty-anchor filter on the reference's training shape: BEV 800x600x27,
30 000 anchors (4 bases x 100x75 positions, stride 8), limit 0, 64 frames per call.

Lines printed (one JSON object each):
  filter_resident   the filter alone on 64 resident views, per-kernel times, occupancy kernel against HBM
  encode_and_filter encode 64 sweeps + filter them, nothing leaves the device between the two
  encode_filter_fused  the same through mv3d_encode_filter_batch_dev (occupancy reported by the tile kernel)
  exact_sums        the float32-sum kernel forced (what a view with negative values or limit != 0 costs)
  host_call         remove_empty_anchor(view, anchors, 0) with numpy buffers (52 MB upload inside, like the reference)
  reference_kernel  the reference's own kernel (oracle/_ref cubin) on this GPU, same call shape, upload included
  cpu_restatement   the C restatement on one host core
Lives under tests/ because its last two legs run the checker (oracle/), which product code and tools/ may not.
python tests/bench_anchors.py [--frames 64] [--steps 20]"""
import argparse
import ctypes
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import numpy as np
import torch

from mv3d_b200.anchors import AnchorFilter, make_anchors
from mv3d_b200.encoder import ViewEncoder
from mv3d_b200.params import REFERENCE_KITTI_TOP as TOP
from mv3d_b200.sweeps import hdl64_sweep, pack_sweeps

ap = argparse.ArgumentParser()
ap.add_argument("--frames", type=int, default=64)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--skip-cpu", action="store_true")
args = ap.parse_args()
B, STEPS = args.frames, args.steps

peaks = {}
try:
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
except Exception:
    pass
HBM = float(peaks.get("hbm_gbs", 0) or 0) or 6500.0

pts, offs = pack_sweeps([hdl64_sweep(f) for f in range(B)])
d_pts = torch.from_numpy(pts).cuda()
enc = ViewEncoder(0, semantics="top")
boxes = make_anchors(TOP.shape[:2], stride=8)
filt = AnchorFilter(enc, anchors=boxes)
top = torch.empty((B, *TOP.shape), dtype=torch.float32, device="cuda")
enc.encode_device(d_pts, offs, TOP, top_out=top)
torch.cuda.synchronize()
lib, ctx = enc._lib, enc._ctx
view_bytes = 4 * int(np.prod(TOP.shape))


def kernel_times(reset=True):
    n = lib.mv3d_kernel_count()
    kms, cnt = (ctypes.c_double * n)(), (ctypes.c_longlong * n)()
    lib.mv3d_ctx_kernel_times(ctx, kms, cnt, int(reset))
    return {lib.mv3d_kernel_name(k).decode(): kms[k] for k in range(n) if cnt[k]}


def timed(fn, steps=STEPS, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    kernel_times()
    lib.mv3d_ctx_set_profiling(ctx, 1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    lib.mv3d_ctx_set_profiling(ctx, 0)
    ms = e0.elapsed_time(e1) / steps
    return ms, {k: round(v / steps, 4) for k, v in kernel_times().items()}


keep = count = None


def run_filter():
    global keep, count
    keep, count = filt.filter_device(top, 0.0)


ms, k = timed(run_filter)
occ_ms = k.get("anchor_occupancy_kernel", 0.0)
print(json.dumps({"case": "filter_resident", "frames": B, "anchors": int(boxes.shape[0]), "ms_per_call": round(ms, 4),
                  "frames_per_s": round(B / (ms * 1e-3), 1), "kernels_ms": k,
                  "kept_mean": float(count.float().mean()),
                  "occupancy_GBps": round(B * view_bytes / (occ_ms * 1e-3) / 1e9, 1) if occ_ms else None,
                  "hbm_peak_GBps": HBM, "occupancy_frac": round(B * view_bytes / (occ_ms * 1e-3) / 1e9 / HBM, 3) if occ_ms else None}))


def run_both():
    enc.encode_device(d_pts, offs, TOP, top_out=top)
    run_filter()


ms2, k2 = timed(run_both)
print(json.dumps({"case": "encode_and_filter", "frames": B, "ms_per_call": round(ms2, 4),
                  "sweeps_per_s": round(B / (ms2 * 1e-3), 1), "kernels_ms": k2}))

def run_fused():
    global keep, count
    _, _, keep, count = filt.encode_and_filter(d_pts, offs, TOP, top_out=top)


ms4, k4 = timed(run_fused)
print(json.dumps({"case": "encode_filter_fused", "frames": B, "ms_per_call": round(ms4, 4),
                  "sweeps_per_s": round(B / (ms4 * 1e-3), 1), "kernels_ms": k4,
                  "kept_mean": float(count.float().mean())}))

exact = AnchorFilter(enc, mode="exact")
exact.n = filt.n


def run_exact():
    exact.filter_device(top[:8], 0.0)


ms3, k3 = timed(run_exact, steps=5, warmup=1)
print(json.dumps({"case": "exact_sums", "frames": 8, "ms_per_call": round(ms3, 4), "ms_per_frame": round(ms3 / 8, 4),
                  "kernels_ms": k3}))

# the reference's call shape: numpy view + numpy anchors in, indices out
view_host = top[0].cpu().numpy()
torch.cuda.synchronize()
host = AnchorFilter(enc)
for _ in range(2):
    got = host.filter_host(view_host, boxes, 0.0)
t0 = time.perf_counter()
for _ in range(10):
    got = host.filter_host(view_host, boxes, 0.0)
host_ms = (time.perf_counter() - t0) / 10 * 1e3
print(json.dumps({"case": "host_call", "ms_per_frame": round(host_ms, 3), "h2d_bytes": view_bytes, "kept": int(len(got))}))

try:
    import ref_anchor_kernel as ref
    if ref.available():
        ref.channel_sums(view_host, boxes[:64])
        t0 = time.perf_counter()
        for _ in range(5):
            inds = ref.channel_sums(view_host, boxes)
            idx = np.where(np.sum(inds, axis=1) > 0)[0]
        ref_ms = (time.perf_counter() - t0) / 5 * 1e3
        print(json.dumps({"case": "reference_kernel", "ms_per_frame": round(ref_ms, 3), "kept": int(len(idx)),
                          "note": "reference .cu compiled for sm_100a, launched as remove_empty_box.py does; racy"}))
except Exception as e:  # the cubin is test infrastructure: its absence is not a bench failure
    print(json.dumps({"case": "reference_kernel", "unavailable": str(e)[:200]}))

if not args.skip_cpu:
    from oracle import anchors as oracle_anchors
    oracle_anchors.build()
    t0 = time.perf_counter()
    want = oracle_anchors.remove_empty_anchor(view_host, boxes, 0.0)
    cpu_ms = (time.perf_counter() - t0) * 1e3
    print(json.dumps({"case": "cpu_restatement", "ms_per_frame": round(cpu_ms, 2), "cores": 1, "kept": int(len(want)),
                      "equal_to_gpu": bool(np.array_equal(want, got))}))
