#!/usr/bin/env python
"""One launch of every kernel of the three reduction strategies, for ncu (run on the GPU box):

  python tools/strategy_profile.py                      # plain run first (must exit 0)
  ncu --set full --clock-control none --import-source on --profile-from-start off \
      -o gpurun_out/strategies_full python tools/strategy_profile.py

Config 2 (64 KITTI-shaped sweeps, BEV 704x800x27 + FV) and config 5 (8 clouds of 2 M points, BEV 1408x1600x10), each
through TILE, ATOMIC and SORT. Every (config, strategy) pair is warmed up once outside the profiled range, then run
once between cudaProfilerStart/Stop, so the report holds exactly one launch of each kernel per pair, in this order.
profiles/summarize_ncu.py turns the report into the table DESIGN.md 5.1 cites (DRAM bytes, L2 atomic / reduction
sectors, shared-memory bank conflicts, achieved GB/s).
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

from mv3d_b200.encoder import ViewEncoder  # noqa: E402
from mv3d_b200.params import DENSE_TOP, KITTI_FRONT_64x512, KITTI_PAPER_TOP  # noqa: E402
from mv3d_b200.sweeps import dense_cloud, hdl64_sweep, pack_sweeps  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    strategies = sys.argv[1:] or ["tile", "atomic", "sort"]
    cases = []
    pts, offs = pack_sweeps([hdl64_sweep(i) for i in range(64)])
    cases.append(("config2", torch.from_numpy(pts).to(dev), offs, KITTI_PAPER_TOP, KITTI_FRONT_64x512, "topfront"))
    cloud = dense_cloud()
    pts, offs = pack_sweeps([cloud] * 8)
    cases.append(("config5", torch.from_numpy(pts).to(dev), offs, DENSE_TOP, None, "top"))
    for name, d, offs, top_p, front_p, sem in cases:
        nb = len(offs) - 1
        top = torch.empty((nb, *top_p.shape), dtype=torch.float32, device=dev)
        front = torch.empty((nb, *front_p.shape), dtype=torch.float32, device=dev) if front_p is not None else None
        for strategy in strategies:
            enc = ViewEncoder(0, semantics=sem, strategy=strategy)
            for rep in range(2):
                enc.encode_device(d, offs, top_p, front_p, top_out=top, front_out=front)
            torch.cuda.synchronize()
            enc.kernel_times(reset=True)
            enc.set_profiling(True)
            torch.cuda.profiler.start()
            enc.encode_device(d, offs, top_p, front_p, top_out=top, front_out=front)
            torch.cuda.synchronize()
            torch.cuda.profiler.stop()
            enc.set_profiling(False)
            kt = enc.kernel_times(reset=True)
            print(name, strategy, {k: round(v[0], 4) for k, v in kt.items() if v[1]}, flush=True)
            enc.close()
        del top, front
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
