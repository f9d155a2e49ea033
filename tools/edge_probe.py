import sys, numpy as np, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import oracle; oracle.build()
from mv3d_b200.encoder import ViewEncoder
from mv3d_b200.params import TopParams, FrontParams
from mv3d_b200.sweeps import uniform_in_box, pack_sweeps
from mv3d_b200.ver3 import createTopAndFrontMaps, createTopMaps
FR = FrontParams(-20.0, 20.0, -45.0, 45.0, 2.5, 5.0, 16, 18, 3)
bad = 0
for name, top in [("deep40000", TopParams(0.0, 3.0, 0.0, 3.0, 0.0, 40.0, 1.0, 1.0, 0.001, 3, 3, 40000)),
                  ("deep500", TopParams(0.0, 6.0, 0.0, 5.0, 0.0, 5.0, 0.5, 0.5, 0.01, 12, 10, 500)),
                  ("wide1row", TopParams(0.0, 0.2, 0.0, 700.0, -1.0, 1.0, 0.1, 0.01, 0.5, 2, 70000, 4)),
                  ("zn0", TopParams(0.0, 4.0, 0.0, 4.0, 0.0, 1.0, 0.5, 0.5, 1.0, 8, 8, 0)),
                  ("one_cell", TopParams(0.0, 1.0, 0.0, 1.0, 0.0, 1.0, 1.0, 1.0, 1.0, 1, 1, 1))]:
    pts = uniform_in_box(20000, top, seed=3, margin=0.2)
    sweeps = [pts, pts[:7], pts[::-1].copy()]
    packed, offs = pack_sweeps(sweeps)
    d = torch.from_numpy(packed).cuda()
    want = [oracle.top_and_front_maps(s, top, FR) for s in sweeps]
    for strategy in ("auto", "tile", "atomic", "sort", "patch"):
        try:
            enc = ViewEncoder(0, strategy=strategy)
            t, f = enc.encode_device(d, offs, top, FR)
            torch.cuda.synchronize()
            ok = all(np.array_equal(t[i].cpu().numpy(), want[i][0], equal_nan=True) and np.array_equal(f[i].cpu().numpy(), want[i][1]) for i in range(3))
            res = "ok" if ok else "MISMATCH"
            enc.close()
        except Exception as e:
            res = "error: " + str(e)[:90]
        if res != "ok": bad += 1
        print(name, strategy, enc.resolved_strategy(top) if 'enc' in dir() else '?', res)
    out = np.ones(top.shape, np.float32); fo = np.ones(FR.shape, np.float32)
    try:
        createTopAndFrontMaps(pts, len(pts), out, fo, top.as_top_paras(), FR.as_front_paras())
        ok = np.array_equal(out, want[0][0], equal_nan=True) and np.array_equal(fo, want[0][1])
        print(name, "host symbol", "ok" if ok else "MISMATCH"); bad += (not ok)
    except Exception as e:
        print(name, "host symbol error", str(e)[:100]); bad += 1
print("bad =", bad)
