"""How the reference calls the path (threading row): from several threads of one process (ctypes
releases the GIL around the call) and from `multiprocessing.Pool` workers forked before the first CUDA use."""
import os
import subprocess
import sys
import threading

import numpy as np
import pytest

from mv3d_b200.params import KITTI_FRONT_64x512, KITTI_PAPER_TOP
from mv3d_b200.sweeps import hdl64_sweep, pack_sweeps

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


def test_reference_symbols_from_concurrent_threads(oracle_mod):
    """Eight threads hammer the two reference symbols on the process-wide default context at once."""
    from mv3d_b200.ver3 import createTopAndFrontMaps, createTopMaps
    top_p, front_p = KITTI_PAPER_TOP, KITTI_FRONT_64x512
    sweeps = [hdl64_sweep(60 + i)[: 120000 - 7919 * i] for i in range(8)]
    errors = []

    def work(i):
        try:
            s = sweeps[i]
            for rep in range(3):
                top = np.full(top_p.shape, float(rep), np.float32)
                if (i + rep) % 2:
                    front = np.full(front_p.shape, 3.0, np.float32)
                    createTopAndFrontMaps(s, s.shape[0], top, front, top_p.as_top_paras(), front_p.as_front_paras())
                    et, ef = oracle_mod.top_and_front_maps(s, top_p, front_p)
                    if not (np.array_equal(top, et) and np.array_equal(front, ef)):
                        errors.append(f"thread {i} rep {rep}: createTopAndFrontMaps differs")
                else:
                    createTopMaps(s, s.shape[0], top, top_p.as_top_paras())
                    if not np.array_equal(top, oracle_mod.top_maps(s, top_p)):
                        errors.append(f"thread {i} rep {rep}: createTopMaps differs")
        except Exception as e:  # noqa: BLE001
            errors.append(f"thread {i}: {e!r}")

    threads = [threading.Thread(target=work, args=(i,)) for i in range(8)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors


def test_one_encoder_from_concurrent_threads(oracle_mod):
    """One context, four threads, each with its own CUDA stream and its own batch: calls are serialised by the context's
    mutex and must not mix their workspaces."""
    import torch
    from mv3d_b200.encoder import ViewEncoder
    top_p, front_p = KITTI_PAPER_TOP, KITTI_FRONT_64x512
    enc = ViewEncoder(0)
    batches = [[hdl64_sweep(70 + 6 * i + j)[: 100000 + 4000 * j] for j in range(n)] for i, n in enumerate((1, 2, 4, 6))]  # both AUTO branches
    out, errors = [None] * 4, []

    def work(i):
        try:
            torch.cuda.set_device(0)
            pts, offs = pack_sweeps(batches[i])
            stream = torch.cuda.Stream()
            with torch.cuda.stream(stream):
                d = torch.from_numpy(pts).cuda()
                for _ in range(3):
                    top, front = enc.encode_device(d, offs, top_p, front_p, stream=stream.cuda_stream)
                stream.synchronize()
            out[i] = (top.cpu().numpy(), front.cpu().numpy())
        except Exception as e:  # noqa: BLE001
            errors.append(f"thread {i}: {e!r}")

    threads = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for i, sweeps in enumerate(batches):
        for j, s in enumerate(sweeps):
            et, ef = oracle_mod.top_and_front_maps(s, top_p, front_p)
            assert np.array_equal(out[i][0][j], et), f"thread {i} sweep {j}: top differs"
            assert np.array_equal(out[i][1][j], ef), f"thread {i} sweep {j}: front differs"
    enc.close()


def test_multiprocessing_pool_fork_then_first_use():
    """The reference's Pool fan-out (src/data.py:495-506): fork first, create the CUDA context in the worker."""
    r = subprocess.run([sys.executable, os.path.join(HERE, "fork_pool_case.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "fork pool ok" in r.stdout
