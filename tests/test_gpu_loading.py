"""GPU tests of the callers either side of the encoder: the KittiLoading-shaped loader and the preprocessing run
(src/utils/batch_loading.py:550-798, src/preprocess_for_mv3d.py) against the Python-twin oracle and numpy's files."""
import io
import os

import numpy as np
import pytest

from mv3d_b200.params import KITTI_FRONT_64x512, PYTWIN_KITTI, REFERENCE_KITTI_TOP
from mv3d_b200.sweeps import hdl64_sweep

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pytwin():
    from oracle import pytwin
    pytwin.build()
    return pytwin


@pytest.fixture(scope="module")
def kitti_tree(tmp_path_factory):
    root = tmp_path_factory.mktemp("kitti")
    for d in ("velodyne", "label_2"):
        os.makedirs(root / "training" / d)
    sweeps = {}
    for i in range(5):
        tag = f"{i:06d}"
        pts = hdl64_sweep(i)[:: 5 + i].copy()
        pts.tofile(root / "training" / "velodyne" / (tag + ".bin"))
        (root / "training" / "label_2" / (tag + ".txt")).write_text("Car 0.0 0 -1.5 1 2 3 4 1.5 1.6 3.9 1 1.7 20 0\n")
        sweeps[tag] = pts
    return str(root), sweeps


@pytest.mark.parametrize("output", ["numpy", "torch"])
def test_loader_python_twin(kitti_tree, pytwin, output):
    from mv3d_b200.loading import KittiLoading
    root, sweeps = kitti_tree
    with KittiLoading(root, is_testset=True, batch_size=2, params=PYTWIN_KITTI, output=output) as loader:
        seen = 0
        for tag, rgb, raw_lidar, top, front in loader:
            if output == "torch":
                assert top.is_cuda and front.is_cuda
                top, front = top.cpu().numpy(), front.cpu().numpy()
            for k, t in enumerate(tag):
                assert np.array_equal(raw_lidar[k], sweeps[t])
                assert np.array_equal(top[k], pytwin.lidar_to_top(sweeps[t], PYTWIN_KITTI)), t
                assert np.array_equal(front[k], pytwin.lidar_to_front_fast(sweeps[t], PYTWIN_KITTI)), t
                seen += 1
        assert seen == 5


def test_loader_c_routine_training(kitti_tree, oracle_mod):
    from mv3d_b200.loading import KittiLoading
    root, sweeps = kitti_tree
    loader = KittiLoading(root, is_testset=False, batch_size=3, params=REFERENCE_KITTI_TOP, front=KITTI_FRONT_64x512)
    tag, label, rgb, raw_lidar, top, front = loader.load()
    assert label[0][0].startswith("Car")
    for k, t in enumerate(tag):
        et, ef = oracle_mod.top_and_front_maps(sweeps[t], REFERENCE_KITTI_TOP, KITTI_FRONT_64x512)
        assert np.array_equal(top[k], et) and np.array_equal(front[k], ef)
    loader.close()


def test_preprocess_run_writes_numpys_files(kitti_tree, pytwin, tmp_path):
    """The files of src/preprocess_for_mv3d.py:36-37: np.save(top_view/<tag>.npy, lidar_to_top(scan)) etc."""
    from mv3d_b200.loading import KittiLoading
    from mv3d_b200.preprocess import preprocess_files
    root, sweeps = kitti_tree
    files = sorted(os.path.join(root, "training", "velodyne", t + ".bin") for t in sweeps)
    top_dir, front_dir = str(tmp_path / "top_view"), str(tmp_path / "front_view")
    stats = preprocess_files(files, top_dir, front_dir, PYTWIN_KITTI, batch=2)
    assert stats["frames"] == 5 and stats["points"] == sum(s.shape[0] for s in sweeps.values())
    for t, s in sweeps.items():
        for d, want in ((top_dir, pytwin.lidar_to_top(s, PYTWIN_KITTI)),
                        (front_dir, pytwin.lidar_to_front_fast(s, PYTWIN_KITTI))):
            buf = io.BytesIO()
            np.save(buf, want)
            assert open(os.path.join(d, t + ".npy"), "rb").read() == buf.getvalue(), (d, t)
    # resume (src/data.py:497-500, overwrite=False): only the frames with a missing output are encoded again
    tags = sorted(sweeps)
    os.unlink(os.path.join(front_dir, tags[1] + ".npy"))
    os.unlink(os.path.join(top_dir, tags[3] + ".npy"))
    marker = os.path.join(top_dir, tags[0] + ".npy")
    open(marker, "wb").write(b"left alone")
    stats = preprocess_files(files, top_dir, front_dir, PYTWIN_KITTI, batch=2, overwrite=False)
    assert stats["skipped"] == 3 and stats["frames"] == 2
    assert open(marker, "rb").read() == b"left alone"
    for t in (tags[1], tags[3]):
        buf = io.BytesIO()
        np.save(buf, pytwin.lidar_to_top(sweeps[t], PYTWIN_KITTI))
        assert open(os.path.join(top_dir, t + ".npy"), "rb").read() == buf.getvalue()
        assert os.path.isfile(os.path.join(front_dir, t + ".npy"))
    preprocess_files([files[0]], top_dir, None, PYTWIN_KITTI)  # default overwrite=True: rewritten
    assert open(marker, "rb").read() != b"left alone"
    # compressed archives (src/data.py:521) load back to the same tensors
    zdir = str(tmp_path / "top")
    preprocess_files(files[:2], zdir, None, PYTWIN_KITTI, batch=2, compressed_top=True)
    for t in list(sweeps)[:2]:
        got = np.load(os.path.join(zdir, t + ".npy.npz"))["top_view"]      # src/utils/batch_loading.py:51
        assert np.array_equal(got, pytwin.lidar_to_top(sweeps[t], PYTWIN_KITTI))
    # and the loader picks the precomputed views up (use_precal_view), never touching the encoder
    os.symlink(top_dir, os.path.join(root, "training", "top_view"))
    os.symlink(front_dir, os.path.join(root, "training", "front_view"))
    try:
        def refuse(*a, **k):
            raise AssertionError("the views are on disk: nothing to encode")
        loader = KittiLoading(root, is_testset=True, batch_size=5, use_precal_view=True, params=PYTWIN_KITTI,
                              encode=refuse)
        tag, rgb, raw_lidar, top, front = loader.load()
        assert np.array_equal(top[3], pytwin.lidar_to_top(sweeps[tag[3]], PYTWIN_KITTI))
        loader.close()
    finally:
        os.unlink(os.path.join(root, "training", "top_view"))
        os.unlink(os.path.join(root, "training", "front_view"))
