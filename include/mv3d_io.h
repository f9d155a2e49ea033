/*
 * mv3d_io.h -- C ABI of the on-disk formats either side of the view encoder.
 * Same library as mv3d_encode.h (libmv3d_encode.so). Host code only: no GPU is needed for these calls; the buffers
 * they fill or drain are the (pinned) host buffers the encoder's host entry points take and return.
 *
 * Formats, as the reference reads and writes them:
 *   sweeps   KITTI velodyne `.bin`: raw little-endian float32 (x, y, z, intensity) records;
 *            `np.fromfile(path, dtype=np.float32).reshape((-1, 4))`
 *            (src/utils/batch_loading.py:656, src/raw_data.py:137-139, src/preprocess_for_mv3d.py:30)
 *   views    `np.save(path, view)` -> `.npy` version 1.0 (src/preprocess_for_mv3d.py:36-37,56-57), read back with
 *            `np.load` (src/utils/batch_loading.py:659-667);
 *            `np.savez_compressed(path, top_view=top)` -> a zip archive holding one deflated `top_view.npy`
 *            (src/data.py:521), read back with `np.load(...)['top_view']` (src/utils/batch_loading.py:51)
 * Every function returns 0 on success, -1 on failure with the reason in mv3d_last_error().
 */
#ifndef MV3D_IO_H
#define MV3D_IO_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MV3D_NPY_MAX_DIMS 8

/* Points per sweep file (file size / 16). A size that is not a multiple of 16 bytes is an error, as it is for the
 * reference's reshape((-1, 4)). */
int mv3d_io_stat_sweeps(const char *const *paths, int n, long long *points_out);

/* Read n sweep files into one packed buffer: sweep i goes to dst + 4*offsets[i] and must hold exactly
 * offsets[i+1] - offsets[i] points (the sizes mv3d_io_stat_sweeps reported). `threads` readers work in parallel
 * (<= 0: one per file, at most 16). dst is typically page-locked (mv3d_host_alloc), so the upload that follows is
 * a direct DMA. */
int mv3d_io_read_sweeps(const char *const *paths, int n, float *dst, const long long *offsets, int threads);

/* The same packing for sweeps that are already in host memory (a ROS subscriber, an earlier np.fromfile): sweep i
 * (offsets[i+1] - offsets[i] points at sweeps[i]) is copied to dst + 4*offsets[i] by `threads` threads in parallel. */
int mv3d_io_pack_sweeps(const float *const *sweeps, int n, float *dst, const long long *offsets, int threads);

/* np.save(path, array) for a C-contiguous float32 array: byte-identical to numpy's file (format 1.0, the header
 * dictionary padded to a multiple of 64 bytes as numpy >= 1.14 pads it). `path` is used as given (numpy appends
 * ".npy" when it is missing; the Python mirror does the same before calling). */
int mv3d_io_write_npy(const char *path, const float *data, const long long *shape, int ndim);

/* The same for n arrays of one shape, array i at data + i*stride_floats, written by `threads` writers in parallel
 * (<= 0: min(n, 16)): the per-frame np.save loop of src/preprocess_for_mv3d.py over a batch the encoder returned. */
int mv3d_io_write_npy_batch(const char *const *paths, int n, const float *data, size_t stride_floats,
                            const long long *shape, int ndim, int threads);

/* Parse the header of a float32, C-order `.npy` file (format 1.0 / 2.0 / 3.0): shape, and the byte offset at which
 * the data starts. Other dtypes or Fortran order are an error. */
int mv3d_io_npy_info(const char *path, long long *shape_out, int *ndim_out, long long *data_offset_out);

/* np.load(path) of such a file into dst (room for capacity_floats; the element count must fit). */
int mv3d_io_read_npy(const char *path, float *dst, size_t capacity_floats);

/* np.savez_compressed(path, <member>=array) for one float32 array: a zip archive with the single deflated member
 * "<member>.npy" (zlib level `level`, numpy uses 6; <= 0 selects 6). np.load reads it back; the archive is not
 * byte-identical to numpy's (zip timestamps; numpy also writes zip64 extra fields). */
int mv3d_io_write_npz_compressed(const char *path, const char *member, const float *data, const long long *shape,
                                 int ndim, int level);
/* n archives of one shape, compressed by `threads` workers in parallel (<= 0: min(n, 16)). */
int mv3d_io_write_npz_compressed_batch(const char *const *paths, int n, const char *member, const float *data,
                                       size_t stride_floats, const long long *shape, int ndim, int level, int threads);

#ifdef __cplusplus
}
#endif
#endif /* MV3D_IO_H */
