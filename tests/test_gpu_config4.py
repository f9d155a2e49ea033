"""GPU test of BASELINE.json configs[3] ("config 4"): a stream of sweeps whose inputs are pre-staged in HBM, encoded
into a ring of output slots, with the per-frame checksum kernel as the consumer -- nothing leaves the device but
16 bytes per frame. Every frame's checksums must equal those of the ORACLE's tensors; the reference idiom replaced is
the frame fan-out of src/data.py:495-506 / src/preprocess_for_mv3d.py:62-72 (one frame per pool worker).
The multi-GPU form shards frames i -> rank i mod world (tests/test_shard_gloo.py covers the host logic on CPU;
bench.py's `also.config4` leg times it under torchrun)."""
import numpy as np
import pytest

from mv3d_b200.params import KITTI_FRONT_64x512, KITTI_PAPER_TOP
from mv3d_b200.sweeps import hdl64_sweep, pack_sweeps

pytestmark = pytest.mark.gpu


def checksum(view: np.ndarray) -> int:
    return int(np.ascontiguousarray(view).view(np.uint32).astype(np.uint64).sum())


@pytest.mark.parametrize("strategy", ["auto", "atomic"])
def test_config4_resident_stream_checksums(torch_cuda, oracle_mod, strategy):
    """520 frames = 8 full batches + a ragged tail of 8, 40 distinct sweeps (some truncated, one empty), two output
    slots. All 520 frames' (BEV, FV) checksums against the oracle's tensors."""
    torch = torch_cuda
    from mv3d_b200.encoder import ViewEncoder
    from mv3d_b200.stream import ResidentStream
    top_p, front_p = KITTI_PAPER_TOP, KITTI_FRONT_64x512
    distinct, frames, batch = 40, 520, 64

    def sweep(key):
        s = hdl64_sweep(300 + key)
        if key % 7 == 3:
            return s[: 120000 - 11000 * (key % 5)].copy()  # ragged point counts
        if key == 11:
            return s[:0].copy()  # an empty sweep in the stream
        return s

    clouds = [sweep(k) for k in range(distinct)]
    want = []
    for c in clouds:
        et, ef = oracle_mod.top_and_front_maps(c, top_p, front_p)
        want.append((checksum(et), checksum(ef)))

    ids = list(range(frames))
    groups = [ids[i:i + batch] for i in range(0, frames, batch)]
    staged = []
    for g in groups:  # pre-stage every batch in HBM before anything is encoded
        pts, offs = pack_sweeps([clouds[f % distinct] for f in g])
        staged.append((torch.from_numpy(pts).cuda(), offs))
    stream = ResidentStream(0, top_p, front_p, batch=batch, depth=2, strategy=strategy)
    chk = ViewEncoder(0)
    sums = torch.zeros((2, frames), dtype=torch.int64, device="cuda")
    done = 0
    for k, top, front in stream.run(staged):
        nb = top.shape[0]
        chk.frame_checksums(top, out=sums[0, done:done + nb])
        chk.frame_checksums(front, out=sums[1, done:done + nb])
        done += nb
    torch.cuda.synchronize()
    assert done == frames and stream.launch_count > 0
    got = sums.cpu().numpy().astype(np.uint64)
    for f in range(frames):
        assert (int(got[0, f]), int(got[1, f])) == want[f % distinct], f"frame {f} (sweep {f % distinct}): checksum differs from the oracle's"
    # the last slot written still holds the ragged tail: compare it whole
    tail = groups[-1]
    top, front = stream.ring[(len(groups) - 1) % 2]
    for j, f in enumerate(tail[:3]):
        et, ef = oracle_mod.top_and_front_maps(clouds[f % distinct], top_p, front_p)
        assert np.array_equal(top[j].cpu().numpy(), et) and np.array_equal(front[j].cpu().numpy(), ef)
    stream.close()
    chk.close()


@pytest.mark.parametrize("strategy,batch,reps", [("tile", 64, 300), ("atomic", 4, 300)])
def test_back_to_back_launches_are_deterministic(torch_cuda, oracle_mod, strategy, batch, reps):
    """Soak: the same batch encoded hundreds of times back to back with no host synchronisation in between, alternating
    two output slots -- the launches are chained by programmatic dependent launch (each kernel starts while its
    predecessor drains), so a missing dependency would show as a rare wrong frame. Every frame's checksum of every
    repetition must equal the oracle's."""
    torch = torch_cuda
    from mv3d_b200.encoder import ViewEncoder
    top_p, front_p = KITTI_PAPER_TOP, KITTI_FRONT_64x512
    distinct = min(batch, 16)
    clouds = [hdl64_sweep(700 + k)[: 120000 - 7000 * (k % 4)] for k in range(distinct)]
    want = np.array([[checksum(t), checksum(f)] for t, f in (oracle_mod.top_and_front_maps(c, top_p, front_p) for c in clouds)],
                    dtype=np.uint64)
    pts, offs = pack_sweeps([clouds[i % distinct] for i in range(batch)])
    d = torch.from_numpy(pts).cuda()
    enc = ViewEncoder(0, strategy=strategy)
    slots = [(torch.empty((batch, *top_p.shape), device="cuda"), torch.empty((batch, *front_p.shape), device="cuda")) for _ in range(2)]
    sums = torch.zeros((reps, 2, batch), dtype=torch.int64, device="cuda")
    for r in range(reps):
        top, front = slots[r % 2]
        enc.encode_device(d, offs, top_p, front_p, top_out=top, front_out=front)
        enc.frame_checksums(top, out=sums[r, 0])
        enc.frame_checksums(front, out=sums[r, 1])
    torch.cuda.synchronize()
    got = sums.cpu().numpy().astype(np.uint64)
    expect = np.stack([want[np.arange(batch) % distinct, 0], want[np.arange(batch) % distinct, 1]])
    bad = np.argwhere(got != expect[None])
    assert bad.size == 0, f"{strategy}: {len(bad)} (repetition, view, frame) checksums differ, first {bad[:5].tolist()}"
    enc.close()
