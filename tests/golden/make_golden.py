"""Generates tests/golden/ver3_golden.npz by running the REFERENCE's own C code (oracle/_ref, compiled
unmodified from /root/reference/src/lidar_data_preprocess/Python_to_C_Interface/ver3/*.c by oracle/Makefile).

Run in the build container only (the GPU box has no /root/reference):  python tests/golden/make_golden.py
Outputs are stored sparsely (flat indices + values of non-zero cells) to keep the fixture small.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import cases  # noqa: E402


def golden_cases():
    out = cases.quirk_cases() + cases.small_cases()
    out += [c for c in cases.sweep_cases(frames=(0,)) if c[0] in ("hdl64_0", "hdl64_0_zround", "hdl64_0_didi")]
    return out


def pack(store, name, key, arr):
    flat = arr.reshape(-1)
    idx = np.flatnonzero((flat != 0) | np.isnan(flat))
    store[f"{name}/{key}/idx"] = idx.astype(np.int32)
    store[f"{name}/{key}/val"] = flat[idx]


def unpack(store, name, key, shape):
    out = np.zeros(int(np.prod(shape)), dtype=np.float32)
    out[store[f"{name}/{key}/idx"]] = store[f"{name}/{key}/val"]
    return out.reshape(shape)


def main():
    import oracle
    oracle.build()
    assert oracle.reference_available(), "needs /root/reference to build oracle/_ref"
    store = {}
    for name, pts, top, front in golden_cases():
        pack(store, name, "top", oracle.top_maps(pts, top, "reference"))
        t, f = oracle.top_and_front_maps(pts, top, front, "reference")
        pack(store, name, "topfront_top", t)
        pack(store, name, "topfront_front", f)
    path = os.path.join(HERE, "ver3_golden.npz")
    np.savez_compressed(path, **store)
    print("wrote", path, os.path.getsize(path), "bytes,", len(golden_cases()), "cases")


if __name__ == "__main__":
    main()
