import sys, time, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from mv3d_b200.encoder import pinned_empty
from mv3d_b200.params import KITTI_FRONT_64x512 as FRONT, KITTI_PAPER_TOP as TOP
from mv3d_b200.sweeps import hdl64_sweep
from mv3d_b200.ver3 import createTopAndFrontMaps, createTopMaps
s0 = hdl64_sweep(0)
for pinned in (False, True):
    alloc = pinned_empty if pinned else (lambda shape: np.empty(shape, np.float32))
    hp = alloc(s0.shape); hp[...] = s0
    ht, hf = alloc(TOP.shape), alloc(FRONT.shape)
    for _ in range(5):
        createTopAndFrontMaps(hp, hp.shape[0], ht, hf, TOP.as_top_paras(), FRONT.as_front_paras())
    t = []
    for _ in range(50):
        t0 = time.perf_counter()
        createTopAndFrontMaps(hp, hp.shape[0], ht, hf, TOP.as_top_paras(), FRONT.as_front_paras())
        t.append((time.perf_counter() - t0) * 1e3)
    print(json.dumps({"split_first": os.environ.get("MV3D_HOST_SPLIT_FIRST", "1"), "buffers": "page-locked" if pinned else "pageable",
                      "ms_per_call_median": round(float(np.median(t)), 3), "p5": round(float(np.percentile(t, 5)), 3), "p95": round(float(np.percentile(t, 95)), 3)}))
