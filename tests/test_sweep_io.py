"""CPU tests of the on-disk formats (include/mv3d_io.h, mv3d_b200/sweep_io.py) against numpy, which is what the
reference reads and writes them with: `.bin` sweeps == np.fromfile(...).reshape((-1, 4)), `.npy` files byte-identical
to np.save, `.npz` archives that np.load opens like the ones np.savez_compressed writes."""
import io
import os
import zipfile

import numpy as np
import pytest

from mv3d_b200 import sweep_io


def _write_bin(path, n, seed):
    pts = np.random.default_rng(seed).standard_normal((n, 4)).astype(np.float32)
    pts.tofile(path)
    return pts


def test_read_sweeps_equals_fromfile(tmp_path):
    sizes = [1000, 0, 1, 123457, 64]
    paths = [str(tmp_path / f"{i:06d}.bin") for i in range(len(sizes))]
    for i, (p, n) in enumerate(zip(paths, sizes)):
        _write_bin(p, n, i)
    points, offsets = sweep_io.read_sweeps(paths, threads=3, pinned=False)
    assert offsets.dtype == np.int32 and offsets.tolist() == np.concatenate([[0], np.cumsum(sizes)]).tolist()
    for i, p in enumerate(paths):
        want = np.fromfile(p, dtype=np.float32).reshape((-1, 4))  # src/utils/batch_loading.py:656
        assert np.array_equal(points[offsets[i]:offsets[i + 1]], want)
    assert np.array_equal(sweep_io.read_sweep(paths[3]), np.fromfile(paths[3], dtype=np.float32).reshape((-1, 4)))


def test_ragged_or_missing_sweep_is_an_error(tmp_path):
    bad = str(tmp_path / "ragged.bin")
    np.arange(7, dtype=np.float32).tofile(bad)           # 28 bytes: reshape((-1, 4)) raises ValueError
    with pytest.raises(ValueError):
        np.fromfile(bad, dtype=np.float32).reshape((-1, 4))
    with pytest.raises(ValueError, match="whole number"):
        sweep_io.read_sweeps([bad], pinned=False)
    with pytest.raises(ValueError, match="No such file"):
        sweep_io.read_sweeps([str(tmp_path / "absent.bin")], pinned=False)


@pytest.mark.parametrize("shape", [(800, 600, 27), (1500, 100, 3), (5,), (), (0, 4), (3, 1), (123456,),
                                   (10 ** 9 // 10 ** 9, 7, 2), (99999, 2), (1, 1, 1, 1, 2)])
def test_npy_is_byte_identical_to_numpy(tmp_path, shape):
    rng = np.random.default_rng(1)
    view = rng.standard_normal(shape).astype(np.float32) if shape != (800, 600, 27) else \
        np.zeros(shape, np.float32)
    path = str(tmp_path / "view")
    sweep_io.save(path, view)                            # appends .npy like np.save
    buf = io.BytesIO()
    np.save(buf, view)
    got = open(path + ".npy", "rb").read()
    assert got == buf.getvalue()
    back = np.load(path + ".npy")
    assert back.shape == view.shape and np.array_equal(back, view)
    mine = sweep_io.load_view(path + ".npy")
    assert mine.shape == view.shape and np.array_equal(mine, view)


def test_header_lengths_across_the_padding_boundary(tmp_path):
    """First-axis digit counts change the header text; every one must land on numpy's 64-byte padding."""
    for n in (1, 9, 10, 99, 100, 12345, 1234567):
        view = np.zeros((n, 1), np.float32)
        path = str(tmp_path / f"h{n}.npy")
        sweep_io.save(path, view)
        buf = io.BytesIO()
        np.save(buf, view)
        assert open(path, "rb").read() == buf.getvalue(), n


def test_save_batch_and_load_numpy_written_files(tmp_path):
    views = np.random.default_rng(2).random((5, 40, 30, 7), dtype=np.float32)
    paths = [str(tmp_path / "top_view" / f"{i:05d}.npy") for i in range(5)]
    os.makedirs(tmp_path / "top_view")
    sweep_io.save_batch(paths, views, threads=4)
    for i, p in enumerate(paths):
        assert np.array_equal(np.load(p), views[i])
    # files numpy wrote (what the loader finds on disk, src/utils/batch_loading.py:659-663), incl. format 2.0
    np.save(str(tmp_path / "n.npy"), views[2])
    assert np.array_equal(sweep_io.load_view(str(tmp_path / "n.npy")), views[2])
    with open(tmp_path / "v2.npy", "wb") as f:
        np.lib.format.write_array(f, views[3], version=(2, 0))
    assert np.array_equal(sweep_io.load_view(str(tmp_path / "v2.npy")), views[3])
    out = np.empty_like(views[1])
    assert sweep_io.load_view(paths[1], out=out) is out and np.array_equal(out, views[1])


def test_load_view_rejects_what_it_cannot_read(tmp_path):
    np.save(str(tmp_path / "f64.npy"), np.zeros((3, 3)))
    with pytest.raises(OSError, match="float32"):
        sweep_io.load_view(str(tmp_path / "f64.npy"))
    np.save(str(tmp_path / "fortran.npy"), np.asfortranarray(np.zeros((3, 4), np.float32)))
    with pytest.raises(OSError, match="Fortran"):
        sweep_io.load_view(str(tmp_path / "fortran.npy"))
    (tmp_path / "junk.npy").write_bytes(b"not a numpy file at all")
    with pytest.raises(OSError):
        sweep_io.load_view(str(tmp_path / "junk.npy"))
    with pytest.raises(TypeError):
        sweep_io.save(str(tmp_path / "x"), np.zeros((2, 2)))            # float64 views are not written


def test_npz_compressed_round_trip(tmp_path):
    """src/data.py:518-521 writes '<n>.npy' through savez_compressed (file '<n>.npy.npz', member 'top_view');
    src/utils/batch_loading.py:51 reads np.load(file + '.npy.npz')['top_view']."""
    top = np.zeros((200, 150, 27), np.float32)
    rng = np.random.default_rng(3)
    hit = rng.random((200, 150)) < 0.05
    top[hit] = rng.random((int(hit.sum()), 27), dtype=np.float32)
    path = str(tmp_path / "00013.npy")
    sweep_io.savez_compressed(path, top_view=top)
    assert os.path.exists(path + ".npz")
    with np.load(path + ".npz") as z:
        assert z.files == ["top_view"]
        assert np.array_equal(z["top_view"], top)
    with zipfile.ZipFile(path + ".npz") as z:
        assert z.testzip() is None                                       # CRC of the member checks out
        info = z.getinfo("top_view.npy")
        assert info.compress_type == zipfile.ZIP_DEFLATED
        buf = io.BytesIO()
        np.save(buf, top)
        assert z.read("top_view.npy") == buf.getvalue()                  # the member is numpy's .npy, byte for byte
        mine = info.compress_size
    ref = str(tmp_path / "numpy_wrote.npy")
    np.savez_compressed(ref, top_view=top)
    with zipfile.ZipFile(ref + ".npz") as z:
        theirs = z.getinfo("top_view.npy").compress_size
    assert mine == theirs                                                # same zlib, same level, same window


def test_npz_batch(tmp_path):
    views = np.random.default_rng(4).random((4, 30, 20, 5), dtype=np.float32)
    paths = [str(tmp_path / f"{i:05d}.npy") for i in range(4)]
    sweep_io.savez_compressed_batch(paths, views, threads=2)
    for i, p in enumerate(paths):
        assert np.array_equal(np.load(p + ".npz")["top_view"], views[i])
    # archives are reproducible: a fixed timestamp, no zip64 padding
    sweep_io.savez_compressed(str(tmp_path / "again.npy"), top_view=views[0])
    assert open(paths[0] + ".npz", "rb").read() == open(str(tmp_path / "again.npy.npz"), "rb").read()


@pytest.mark.parametrize("threads", [0, 1, 3])
def test_pack_sweeps_into_equals_concatenate(threads):
    """In-memory sweeps packed by the native parallel copier == np.concatenate (ragged, empty and single-point sweeps)."""
    from mv3d_b200.sweep_io import pack_sweeps_into
    rng = np.random.default_rng(5)
    sweeps = [rng.random((n, 4), dtype=np.float32) for n in (1000, 0, 1, 120000, 77, 0, 50001)]
    dst = np.full((sum(len(s) for s in sweeps) + 8, 4), -1.0, np.float32)
    offs = pack_sweeps_into(sweeps, dst, threads)
    assert offs.tolist() == np.concatenate([[0], np.cumsum([len(s) for s in sweeps])]).tolist()
    assert np.array_equal(dst[: offs[-1]], np.concatenate(sweeps))
    assert (dst[offs[-1]:] == -1.0).all()
    assert pack_sweeps_into([], dst).tolist() == [0]
    sliced = rng.random((100, 8), dtype=np.float32)[:, :4]  # not contiguous: copied first
    assert np.array_equal(dst[: pack_sweeps_into([sliced], dst)[-1]], sliced)
