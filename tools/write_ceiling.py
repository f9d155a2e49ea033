#!/usr/bin/env python
"""Write-only and copy bandwidth of this GPU for a buffer the size of one step's output (3.9 GB): the ceiling the
tile kernel's stores can be compared with besides MEASURED_PEAKS.json's copy figure. python tools/write_ceiling.py"""
import json

import torch

n = 64 * 704 * 800 * 27
x = torch.empty(n, dtype=torch.float32, device="cuda")
y = torch.empty(n // 2, dtype=torch.float32, device="cuda")


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


ms_fill = timed(lambda: x.zero_())
ms_copy = timed(lambda: x[: n // 2].copy_(y))
print(json.dumps({"bytes": 4 * n, "memset_ms": round(ms_fill, 4), "memset_GBps": round(4 * n / ms_fill / 1e6, 1),
                  "copy_ms": round(ms_copy, 4), "copy_GBps_read_plus_write": round(4 * n / ms_copy / 1e6, 1)}))
