#!/usr/bin/env python
"""BASELINE.json configs 1, 3 and 5 on one B200 (config 2 is bench.py's line, config 4 tests/run_config4.py), with
the reduction strategies side by side -- SURVEY.md 8(d) asks for the atomic-vs-tile decision to be re-run on the dense
cloud of config 5. Inputs and outputs resident in HBM; CUDA events around the timed launches; one JSON line per case.

  python tools/configs_bench.py [--steps 20] > gpurun_out/configs.jsonl
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from mv3d_b200.encoder import ViewEncoder, pinned_empty
from mv3d_b200.params import DENSE_TOP, KITTI_FRONT_64x512 as FRONT, KITTI_PAPER_TOP as TOP
from mv3d_b200.sweeps import dense_cloud, hdl64_sweep, pack_sweeps
from mv3d_b200.ver3 import createTopAndFrontMaps


def peak():
    try:
        with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"])
    except Exception:
        return 6552.0


def timed(enc, fn, steps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    enc.kernel_times(reset=True)
    enc.set_profiling(True)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    enc.set_profiling(False)
    k = {n: round(v[0] / steps, 4) for n, v in enc.kernel_times(reset=True).items() if v[1]}
    return a.elapsed_time(b) / steps, k


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--crossover", action="store_true", help="only the tile-vs-atomic crossover sweeps")
    args = ap.parse_args()
    pk = peak()

    # ---- config 1: one sweep, device-resident and through the reference symbol with host buffers ----
    s0 = hdl64_sweep(0)
    pts, offs = pack_sweeps([s0])
    d = torch.from_numpy(pts).cuda()
    for strategy in ("tile", "atomic"):
        enc = ViewEncoder(0, semantics="topfront", strategy=strategy)
        top = torch.empty((1, *TOP.shape), dtype=torch.float32, device="cuda")
        front = torch.empty((1, *FRONT.shape), dtype=torch.float32, device="cuda")
        ms, k = timed(enc, lambda: enc.encode_device(d, offs, TOP, FRONT, top_out=top, front_out=front), args.steps * 5)
        by = 16 * len(s0) + 4 * int(np.prod(TOP.shape)) + 4 * int(np.prod(FRONT.shape))
        print(json.dumps({"config": 1, "what": "one 120k-point sweep, BEV 704x800x27 + FV 64x512x3, device-resident",
                          "strategy": strategy, "ms_per_sweep": round(ms, 4), "sweeps_per_s": round(1e3 / ms, 1),
                          "algorithmic_GBps": round(by / ms / 1e6, 1), "frac_of_peak": round(by / ms / 1e6 / pk, 3),
                          "kernels_ms": k}), flush=True)
        enc.close()
    for pinned in (False, True):
        alloc = pinned_empty if pinned else (lambda shape: np.empty(shape, np.float32))
        hp = alloc(s0.shape)
        hp[...] = s0
        ht, hf = alloc(TOP.shape), alloc(FRONT.shape)
        for _ in range(3):
            createTopAndFrontMaps(hp, hp.shape[0], ht, hf, TOP.as_top_paras(), FRONT.as_front_paras())
        t = []
        for _ in range(30):
            t0 = time.perf_counter()
            createTopAndFrontMaps(hp, hp.shape[0], ht, hf, TOP.as_top_paras(), FRONT.as_front_paras())
            t.append((time.perf_counter() - t0) * 1e3)
        print(json.dumps({"config": 1, "what": "the reference symbol createTopAndFrontMaps, host numpy buffers, one call per sweep",
                          "buffers": "page-locked" if pinned else "pageable", "ms_per_call_median": round(float(np.median(t)), 3),
                          "p5": round(float(np.percentile(t, 5)), 3), "p95": round(float(np.percentile(t, 95)), 3)}), flush=True)

    # ---- batch-size crossover of the two strategies on config-1 sweeps (what AUTO decides from) ----
    if args.crossover:
        allsw = [hdl64_sweep(f) for f in range(32)]
        for batch in (1, 2, 3, 4, 6, 8, 16, 32):
            pts, offs = pack_sweeps(allsw[:batch])
            d = torch.from_numpy(pts).cuda()
            top = torch.empty((batch, *TOP.shape), dtype=torch.float32, device="cuda")
            front = torch.empty((batch, *FRONT.shape), dtype=torch.float32, device="cuda")
            row = {"crossover": "config-1 sweeps", "batch": batch}
            for strategy in ("tile", "atomic"):
                enc = ViewEncoder(0, semantics="topfront", strategy=strategy)
                ms, _ = timed(enc, lambda: enc.encode_device(d, offs, TOP, FRONT, top_out=top, front_out=front), args.steps * 2)
                row[strategy + "_ms"] = round(ms, 4)
                enc.close()
            print(json.dumps(row), flush=True)
        for batch in (1, 2, 4, 8):
            pts, offs = pack_sweeps([dense_cloud(frame_id=f) for f in range(batch)])
            d = torch.from_numpy(pts).cuda()
            top = torch.empty((batch, *DENSE_TOP.shape), dtype=torch.float32, device="cuda")
            row = {"crossover": "config-5 clouds", "batch": batch}
            for strategy in ("tile", "atomic"):
                enc = ViewEncoder(0, semantics="top", strategy=strategy)
                ms, _ = timed(enc, lambda: enc.encode_device(d, offs, DENSE_TOP, top_out=top), args.steps)
                row[strategy + "_ms"] = round(ms, 4)
                enc.close()
            print(json.dumps(row), flush=True)
        # thinner and denser clouds on the config-5 grid: where does the point work outweigh the dense write?
        full = dense_cloud(frame_id=0)
        for keep in (8, 4, 2, 1):
            sub = np.ascontiguousarray(full[::keep])
            pts, offs = pack_sweeps([sub] * 8)
            d = torch.from_numpy(pts).cuda()
            top = torch.empty((8, *DENSE_TOP.shape), dtype=torch.float32, device="cuda")
            row = {"crossover": "config-5 grid, 8 clouds, every n-th point", "n": keep, "points_per_cloud": len(sub),
                   "points_per_cell": round(len(sub) / (DENSE_TOP.shape[0] * DENSE_TOP.shape[1]), 3)}
            for strategy in ("tile", "atomic"):
                enc = ViewEncoder(0, semantics="top", strategy=strategy)
                ms, _ = timed(enc, lambda: enc.encode_device(d, offs, DENSE_TOP, top_out=top), args.steps)
                row[strategy + "_ms"] = round(ms, 4)
                enc.close()
            print(json.dumps(row), flush=True)
        return

    # ---- config 3: front view only, 64 sweeps per launch ----
    sweeps = [hdl64_sweep(f) for f in range(64)]
    pts, offs = pack_sweeps(sweeps)
    d = torch.from_numpy(pts).cuda()
    enc = ViewEncoder(0, semantics="topfront")
    front = torch.empty((64, *FRONT.shape), dtype=torch.float32, device="cuda")
    ms, k = timed(enc, lambda: enc.encode_device(d, offs, TOP, FRONT, front_out=front, want_top=False), args.steps)
    by = 16 * len(pts) + 64 * 4 * int(np.prod(FRONT.shape))
    print(json.dumps({"config": 3, "what": "front view only (BEV gate applied, BEV not written), 64 sweeps per launch",
                      "ms_per_64": round(ms, 4), "sweeps_per_s": round(64e3 / ms, 1), "algorithmic_GBps": round(by / ms / 1e6, 1),
                      "frac_of_peak": round(by / ms / 1e6 / pk, 3), "kernels_ms": k}), flush=True)
    enc.close()
    del d, front

    # ---- config 5: dense aggregated cloud, 2 M points, BEV 1408x1600x10, 8 clouds per launch ----
    B = 8
    clouds = [dense_cloud(frame_id=f) for f in range(B)]
    pts, offs = pack_sweeps(clouds)
    d = torch.from_numpy(pts).cuda()
    top = torch.empty((B, *DENSE_TOP.shape), dtype=torch.float32, device="cuda")
    by = 16 * len(pts) + B * 4 * int(np.prod(DENSE_TOP.shape))
    for strategy in ("tile", "atomic", "patch", "sort"):
        enc = ViewEncoder(0, semantics="top", strategy=strategy)
        ms, k = timed(enc, lambda: enc.encode_device(d, offs, DENSE_TOP, top_out=top), args.steps)
        print(json.dumps({"config": 5, "what": "2 M-point cloud, BEV 1408x1600x(8+2) at 0.05 m, 8 clouds per launch",
                          "strategy": strategy, "ms_per_cloud": round(ms / B, 4), "clouds_per_s": round(B * 1e3 / ms, 1),
                          "points_per_s": round(len(pts) * 1e3 / ms), "algorithmic_GBps": round(by / ms / 1e6, 1),
                          "frac_of_peak": round(by / ms / 1e6 / pk, 3), "kernels_ms_per_launch": k}), flush=True)
        enc.close()


if __name__ == "__main__":
    main()
