#!/usr/bin/env python
"""Throughput of the Python-twin mode (device-resident, 64 KITTI-shaped sweeps per launch, reference `kitti` preset:
BEV 800x600x27 + FV 1500x100x3), with per-kernel times. Prints one JSON line.  python tools/pytwin_bench.py"""
import ctypes
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from mv3d_b200.params import PYTWIN_KITTI as P
from mv3d_b200.pytwin import PyTwinEncoder
from mv3d_b200.sweeps import hdl64_sweep, pack_sweeps

BATCH, STEPS = 64, 20
pts, offs = pack_sweeps([hdl64_sweep(f) for f in range(BATCH)])
d = torch.from_numpy(pts).cuda()
enc = PyTwinEncoder(0)
top = torch.empty((BATCH, *P.top_shape), dtype=torch.float32, device="cuda")
front = torch.empty((BATCH, *P.front_shape), dtype=torch.float32, device="cuda")
for _ in range(3):
    enc.encode_device(d, offs, P, top_out=top, front_out=front)
torch.cuda.synchronize()
lib = enc._lib
lib.mv3d_ctx_kernel_times(enc._ctx, None, None, 1)
lib.mv3d_ctx_set_profiling(enc._ctx, 1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(STEPS):
    enc.encode_device(d, offs, P, top_out=top, front_out=front)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / STEPS
n = lib.mv3d_kernel_count()
kms, cnt = (ctypes.c_double * n)(), (ctypes.c_longlong * n)()
lib.mv3d_ctx_kernel_times(enc._ctx, kms, cnt, 1)
kernels = {lib.mv3d_kernel_name(k).decode(): round(kms[k] / STEPS, 4) for k in range(n) if cnt[k]}
bytes_per_sweep = 16 * (len(pts) // BATCH) + 4 * int(np.prod(P.top_shape)) + 4 * int(np.prod(P.front_shape))
print(json.dumps({"mode": "python-twin", "sweeps_per_s": BATCH / (ms * 1e-3), "ms_per_step": ms, "batch": BATCH,
                  "algorithmic_bytes_per_sweep": bytes_per_sweep,
                  "step_GBps": bytes_per_sweep * BATCH / (ms * 1e-3) / 1e9, "kernels_ms_per_step": kernels}))
