"""CPU tests of the KittiLoading-shaped loader's host logic (mv3d_b200/loading.py): file listing, batching, test-set
exhaustion, training wrap-around and shuffle, precomputed views, labels -- with the GPU encode replaced by a stand-in
(the encode itself is covered by the GPU suites)."""
import os

import numpy as np
import pytest

from mv3d_b200.loading import KittiLoading, list_frames
from mv3d_b200.params import PyTwinParams

TINY = PyTwinParams(0, 8, -4, 4, -2, 1, 1.0, 1.0, 1.0, front_width=6, front_height=4)


def make_tree(root, n, with_views=()):
    for d in ("velodyne", "label_2", "image_2", "top_view", "front_view"):
        os.makedirs(os.path.join(root, "training", d), exist_ok=True)
    sweeps = {}
    for i in range(n):
        tag = f"{i:06d}"
        pts = np.random.default_rng(i).standard_normal((50 + 7 * i, 4)).astype(np.float32)
        pts.tofile(os.path.join(root, "training", "velodyne", tag + ".bin"))
        open(os.path.join(root, "training", "label_2", tag + ".txt"), "w").write(f"Car 0 0 0 {i}\nDontCare {i}\n")
        sweeps[tag] = pts
        if i in with_views:
            np.save(os.path.join(root, "training", "top_view", tag + ".npy"),
                    np.full(TINY.top_shape, 100 + i, np.float32))
            np.save(os.path.join(root, "training", "front_view", tag + ".npy"),
                    np.full(TINY.front_shape, 200 + i, np.float32))
    return sweeps


def fake_encode(points, offsets, want_top=True, want_front=True):
    """Stand-in: every view is filled with the sweep's point count."""
    n = len(offsets) - 1
    top = np.stack([np.full(TINY.top_shape, offsets[k + 1] - offsets[k], np.float32) for k in range(n)])
    front = np.stack([np.full(TINY.front_shape, offsets[k + 1] - offsets[k], np.float32) for k in range(n)])
    return top, front


def test_listing_matches_the_reference_layout(tmp_path):
    make_tree(str(tmp_path), 3)
    files = list_frames(str(tmp_path))
    assert files["tag"] == ["000000", "000001", "000002"]
    assert files["lidar"][1].endswith("training/velodyne/000001.bin")
    assert files["top"][2].endswith("training/top_view/000002.npy")
    split = tmp_path / "val.txt"
    split.write_text("000002\n000000\n")
    files = list_frames(str(tmp_path), split_file=str(split))
    assert files["tag"] == ["000002", "000000"]
    assert files["label"][0].endswith("training/label_2/000002.txt")


def test_testset_batches_in_order_then_none(tmp_path):
    sweeps = make_tree(str(tmp_path), 5)
    with KittiLoading(str(tmp_path), is_testset=True, batch_size=2, params=TINY, encode=fake_encode,
                      pinned=False) as loader:
        assert len(loader) == 5 and loader.get_shape()[0] == TINY.top_shape
        seen = []
        while True:
            batch = loader.load()
            if batch is None:
                break
            tag, rgb, raw_lidar, top, front = batch
            assert len(tag) == len(raw_lidar) == top.shape[0] == front.shape[0] == len(rgb)
            for k, t in enumerate(tag):
                assert np.array_equal(raw_lidar[k], sweeps[t])
                assert top[k, 0, 0, 0] == sweeps[t].shape[0] and rgb[k] is None
            seen += list(tag)
            assert loader.get_frame_info() == tag[-1]
        assert seen == sorted(sweeps) and loader.alreay_extract_data == 5
        assert loader.load() is None


def test_training_wraps_around_and_shuffles(tmp_path):
    sweeps = make_tree(str(tmp_path), 4)
    loader = KittiLoading(str(tmp_path), is_testset=False, require_shuffle=True, batch_size=3, params=TINY,
                          encode=fake_encode, seed=7, pinned=False, rgb_reader=lambda p: os.path.basename(p))
    tags = []
    for _ in range(4):                                   # 12 frames out of a 4-frame set: three epochs
        tag, label, rgb, raw_lidar, top, front = loader.load()
        assert label[0][0].startswith("Car") and len(label[0]) == 2
        assert rgb[0] == tag[0] + ".png"
        tags += list(tag)
    loader.close()
    epochs = [tags[i:i + 4] for i in range(0, 12, 4)]
    assert all(sorted(e) == sorted(sweeps) for e in epochs)      # every epoch sees every frame once
    assert len({tuple(e) for e in epochs}) > 1                   # and not always in the same order


def test_precomputed_views_with_fallback(tmp_path):
    make_tree(str(tmp_path), 4, with_views=(0, 2))
    loader = KittiLoading(str(tmp_path), is_testset=True, batch_size=4, use_precal_view=True, params=TINY,
                          encode=fake_encode, pinned=False)
    tag, rgb, raw_lidar, top, front = loader.load()
    assert top[0, 0, 0, 0] == 100 and front[0, 0, 0, 0] == 200          # read from disk
    assert top[2, 0, 0, 0] == 102 and front[2, 0, 0, 0] == 202
    assert top[1, 0, 0, 0] == raw_lidar[1].shape[0]                      # no file: encoded
    assert front[3, 0, 0, 0] == raw_lidar[3].shape[0]
    loader.close()


def test_errors_surface_on_load(tmp_path):
    make_tree(str(tmp_path), 2)
    np.arange(7, dtype=np.float32).tofile(os.path.join(str(tmp_path), "training", "velodyne", "000001.bin"))
    loader = KittiLoading(str(tmp_path), is_testset=True, batch_size=2, params=TINY, encode=fake_encode, pinned=False)
    with pytest.raises(ValueError, match="whole number"):
        loader.load()
    loader.close()


def test_iteration(tmp_path):
    make_tree(str(tmp_path), 3)
    loader = KittiLoading(str(tmp_path), is_testset=True, batch_size=2, params=TINY, encode=fake_encode, pinned=False)
    assert [len(b[0]) for b in loader] == [2, 1]
    loader.close()


def test_preprocess_resume_skips_finished_frames(tmp_path):
    """overwrite=False (the reference's resume, src/data.py:497-500): frames whose outputs exist are neither read nor
    encoded; with every output present the encoder is never touched (so this runs without a GPU)."""
    from mv3d_b200.params import PYTWIN_KITTI
    from mv3d_b200.preprocess import preprocess_files

    class Refuse:
        def encode_host(self, *a, **k):
            raise AssertionError("nothing to encode")

    files = [str(tmp_path / "velodyne" / f"{i:06d}.bin") for i in range(4)]  # never opened: they do not even exist
    top_dir, front_dir, zdir = tmp_path / "top_view", tmp_path / "front_view", tmp_path / "top"
    for d in (top_dir, front_dir, zdir):
        d.mkdir()
    for i in range(4):
        (top_dir / f"{i:06d}.npy").write_bytes(b"x")
        (front_dir / f"{i:06d}.npy").write_bytes(b"x")
        (zdir / f"{i:06d}.npy.npz").write_bytes(b"x")
    stats = preprocess_files(files, str(top_dir), str(front_dir), PYTWIN_KITTI, overwrite=False, encoder=Refuse())
    assert stats["skipped"] == 4 and stats["frames"] == 0
    stats = preprocess_files(files, str(zdir), None, PYTWIN_KITTI, compressed_top=True, overwrite=False, encoder=Refuse())
    assert stats["skipped"] == 4
    (front_dir / "000002.npy").unlink()  # one output missing: that frame is due again, and only that one
    try:
        preprocess_files(files, str(top_dir), str(front_dir), PYTWIN_KITTI, overwrite=False, encoder=Refuse())
    except Exception as e:
        assert "000002" in str(e) or isinstance(e, (OSError, RuntimeError, AssertionError))
    else:
        raise AssertionError("frame 000002 should have been read")
