#!/usr/bin/env python
"""Profiler target: a 64-sweep encode where every sweep has one point (the tile kernel does nothing but bulk stores),
then the same with full sweeps. python tools/one_point_encode.py [reps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from mv3d_b200.encoder import ViewEncoder
from mv3d_b200.params import KITTI_PAPER_TOP as TOP
from mv3d_b200.sweeps import hdl64_sweep, pack_sweeps

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
enc = ViewEncoder(0, semantics="top", strategy="tile")
top = torch.empty((64, *TOP.shape), dtype=torch.float32, device="cuda")
full = [hdl64_sweep(f) for f in range(64)]
for sweeps in ([s[:1] for s in full], full):
    pts, offs = pack_sweeps(sweeps)
    d = torch.from_numpy(pts).cuda()
    for _ in range(reps):
        enc.encode_device(d, offs, TOP, top_out=top)
    torch.cuda.synchronize()
