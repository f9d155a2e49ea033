#!/usr/bin/env python
"""Where the tile kernel's time goes: the same 64-sweep launch on (a) sweeps of one point each -- every tile but one
per sweep is empty, so the kernel is nothing but bulk stores of the zero image -- and (b) HDL-64 sweeps thinned to
1/8, 1/4, 1/2 and all of their points. python tools/tile_ceiling.py"""
import ctypes
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from mv3d_b200.encoder import ViewEncoder
from mv3d_b200.params import KITTI_FRONT_64x512 as FRONT, KITTI_PAPER_TOP as TOP
from mv3d_b200.sweeps import hdl64_sweep, pack_sweeps

B, STEPS = 64, 20
enc = ViewEncoder(0, semantics="topfront", strategy="tile")
lib, ctx = enc._lib, enc._ctx
top = torch.empty((B, *TOP.shape), dtype=torch.float32, device="cuda")
front = torch.empty((B, *FRONT.shape), dtype=torch.float32, device="cuda")
out_bytes = 4 * B * int(np.prod(TOP.shape))
full = [hdl64_sweep(f) for f in range(B)]
for name, sweeps in [("one_point", [s[:1] for s in full]), ("1/8", [s[::8] for s in full]), ("1/4", [s[::4] for s in full]),
                     ("1/2", [s[::2] for s in full]), ("all", full)]:
    pts, offs = pack_sweeps(sweeps)
    d = torch.from_numpy(pts).cuda()
    for _ in range(3):
        enc.encode_device(d, offs, TOP, FRONT, top_out=top, front_out=front)
    torch.cuda.synchronize()
    lib.mv3d_ctx_kernel_times(ctx, None, None, 1)
    lib.mv3d_ctx_set_profiling(ctx, 1)
    for _ in range(STEPS):
        enc.encode_device(d, offs, TOP, FRONT, top_out=top, front_out=front)
    torch.cuda.synchronize()
    lib.mv3d_ctx_set_profiling(ctx, 0)
    n = lib.mv3d_kernel_count()
    kms, cnt = (ctypes.c_double * n)(), (ctypes.c_longlong * n)()
    lib.mv3d_ctx_kernel_times(ctx, kms, cnt, 1)
    k = {lib.mv3d_kernel_name(i).decode(): kms[i] / STEPS for i in range(n) if cnt[i]}
    t = k["tile_build_kernel"]
    print(json.dumps({"sweeps": name, "points_per_sweep": int(len(pts) // B), "tile_build_ms": round(t, 4),
                      "tile_build_GBps": round(out_bytes / t / 1e6, 1),
                      "kernels_ms": {a: round(v, 4) for a, v in k.items()}}))
