/*
 * mv3d_encode.h -- C ABI of libmv3d_encode.so, the B200 (sm_100a) LiDAR view encoder.
 *
 * Part 1 is the drop-in boundary: the two symbols MV3D loads with ctypes.cdll.LoadLibrary, with
 * byte-identical signatures, host pointers, synchronous, same output layout.
 * Part 2 is additive: a device-resident, batched entry point for callers that keep sweeps in HBM.
 *
 * All citations are relative to the reference tree (jeasinema/MV3D).
 * No function here throws or aborts; failures are reported through mv3d_last_error().
 * There is no CPU fallback: without a CUDA device every entry point fails loudly.
 */
#ifndef MV3D_ENCODE_H
#define MV3D_ENCODE_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------------------------------------
 * Part 1 -- reference symbols (drop-in)
 * ---------------------------------------------------------------------------------------------- */

/* Replaces src/lidar_data_preprocess/Python_to_C_Interface/ver3/LidarTopPreprocess.c:30.
 * Bound by src/data.py:280-289 and ver3/createTopMaps.py:15-23.
 *   raw_data : host, num x 4 float32 (x, y, z, intensity), C-contiguous, read-only
 *   top_data : host, X_SIZE*Y_SIZE*(Z_SIZE+2) float32, fully overwritten (despite the const)
 * Cell index by (int) truncation. On failure top_data is zero-filled and mv3d_last_error() is set. */
void createTopMaps(const void *raw_data, int num, const void *top_data,
                   float x_MIN, float x_MAX, float y_MIN, float y_MAX, float z_MIN, float z_MAX,
                   float x_DIVISION, float y_DIVISION, float z_DIVISION,
                   int X_SIZE, int Y_SIZE, int Z_SIZE);

/* Replaces .../ver3/LidarTopAndFrontPreprocess.c:30. Bound by ver3/createTopAndFrontMaps.py:20-33.
 *   front_data : host, R_SIZE*C_SIZE*F_SIZE float32, fully overwritten
 * Cell index by (int)floor(). The reference stores one row past the end of front_data when
 * r_coord == R_SIZE; this implementation applies the same linear index and drops those stores.
 * C_SIZE/R_SIZE/F_SIZE must hold integral values with F_SIZE >= 3 (what the reference's wrapper passes). */
void createTopAndFrontMaps(const void *raw_data, int num, const void *top_data, const void *front_data,
                           float x_MIN, float x_MAX, float y_MIN, float y_MAX, float z_MIN, float z_MAX,
                           float x_DIVISION, float y_DIVISION, float z_DIVISION,
                           int X_SIZE, int Y_SIZE, int Z_SIZE,
                           float phi_MIN, float phi_MAX, float theta_MIN, float theta_MAX,
                           float phi_DIVISION, float theta_DIVISION,
                           float C_SIZE, float R_SIZE, float F_SIZE);

/* ------------------------------------------------------------------------------------------------
 * Part 2 -- additive API (the reference has no equivalent; its ABI returns void and keeps no state)
 * ---------------------------------------------------------------------------------------------- */

/* Error text of the last failing call on this thread ("" if none). Never NULL. */
const char *mv3d_last_error(void);

/* What the reference's Python wrapper does around the C call (src/data.py:265-291 `clidar_to_top`), fused into it:
 *   _FLIP_XY            write cell (X-1, Y-1) at [xn-X][yn-Y]: the result of `np.flipud(np.fliplr(top_flip))` (:290)
 *   _FILTER_CENTER_CAR  drop the points with |x| <= 4.7/2 and |y| <= 2.1/2 first: `filter_center_car` (:224-227),
 *                       applied for the didi data sets (:266-267); the kept points keep their order */
enum { MV3D_TOP_FLIP_XY = 1, MV3D_TOP_FILTER_CENTER_CAR = 2 };

/* The 9 floats + 3 ints of createTopMaps, in the reference's argument order, then MV3D_TOP_* flags (0 = exactly the
 * C routine). */
typedef struct mv3d_top_params {
    float x_min, x_max, y_min, y_max, z_min, z_max;
    float x_div, y_div, z_div;
    int xn, yn, zn;
    int flags;
} mv3d_top_params;

/* The 9 trailing floats of createTopAndFrontMaps, in the reference's argument order. */
typedef struct mv3d_front_params {
    float phi_min, phi_max, theta_min, theta_max; /* dead parameters in the reference, kept for symmetry */
    float phi_div, theta_div;
    float c_size, r_size, f_size;
} mv3d_front_params;

/* Which reference routine's arithmetic to reproduce. They differ in the density channel
 * (truncation keeps points with z in (z_MIN - z_DIV, z_MIN); floor rejects them). */
enum {
    MV3D_SEM_TOP = 0,      /* createTopMaps:          (int)((p - MIN) / DIV)        LidarTopPreprocess.c:107-109 */
    MV3D_SEM_TOPFRONT = 1  /* createTopAndFrontMaps:  (int)floor((p - MIN) / DIV)   LidarTopAndFrontPreprocess.c:124-126 */
};

/* Per-cell reduction strategy for the bird's-eye view. */
enum {
    MV3D_STRATEGY_AUTO = 0,
    MV3D_STRATEGY_ATOMIC = 1, /* zero-fill, then order-preserving integer atomics straight into the output in HBM/L2 */
    MV3D_STRATEGY_TILE = 2,   /* bin points by output tile, reduce in shared memory, one bulk (TMA) store per tile */
    MV3D_STRATEGY_PATCH = 3,  /* zero-fill on a side stream (copy engine) while points are binned; then reduce per tile
                                 in shared memory and overwrite only the touched cells */
    MV3D_STRATEGY_SORT = 4    /* radix sort of the kept points by (frame, cell, slice) key, then one walk per cell run:
                                 the measured alternative to the tile strategy (C-routine semantics, keys must fit 32
                                 bits: batch * Xn * Yn * Zn < 2^32; the sort itself is CUB's, a library kernel) */
};

typedef struct mv3d_ctx mv3d_ctx;

/* Creates a per-GPU context (workspace grows on demand, density table, stream pool). 0 on success.
 * Concurrency: every entry point may be called from any thread (calls on one context are serialised by a mutex).
 * The asynchronous device entry points (`*_dev`, `void *stream` argument) share the context's scratch buffers, so the
 * library orders the work of successive calls on ONE context on the device: a call on a different stream than the
 * previous one first waits (cudaStreamWaitEvent, no host blocking) for that call to finish. Results are correct from
 * any thread and any stream; for work that should really overlap create one context per stream (mv3d_b200/stream.py
 * does exactly that for its ring of slots). */
int mv3d_ctx_create(int device, mv3d_ctx **out);
void mv3d_ctx_destroy(mv3d_ctx *ctx);

/* Encodes `batch` sweeps that are already resident in device memory.
 *   points_dev   : device, total x 4 float32, 16-byte aligned; sweep i is rows [offsets[i], offsets[i+1])
 *   offsets_host : host, batch+1 ints, non-decreasing, offsets[0] >= 0
 *   top_dev      : device, batch x [xn][yn][zn+2] float32, or NULL to skip the bird's-eye view
 *   front_dev    : device, batch x [R][C][F] float32, or NULL to skip the front view (fp may then be NULL)
 *   stream       : a cudaStream_t (NULL = legacy default stream). The call only enqueues work.
 * Returns 0 on success. */
int mv3d_encode_batch_dev(mv3d_ctx *ctx, const float *points_dev, const int *offsets_host, int batch,
                          float *top_dev, float *front_dev,
                          const mv3d_top_params *tp, const mv3d_front_params *fp,
                          int semantics, int strategy, void *stream);

/* Same, with host buffers, pipelined sweep by sweep over internal streams; returns after the outputs are
 * complete in host memory. The bird's-eye view is ~95 % zeros, so by default it is read back COMPACT: the
 * device reports, per touched cell, its index and the finished values of its non-zero planes (about 1 MB per KITTI
 * sweep instead of 60.8 MB) and host threads (MV3D_HOST_THREADS, default min(16, CPUs / LOCAL_WORLD_SIZE); pinned to
 * the rank's share of the CPUs when several ranks share a host, MV3D_HOST_PIN) rebuild the dense tensor in the
 * caller's buffer in one pass of streaming stores -- placement only, no encoding arithmetic on the host. The rate is
 * bound by the host's memory (every sweep's 60.8 MB has to be written there by somebody), not by the GPU.
 * MV3D_HOST_READBACK=dense copies the dense tensor over PCIe instead (pinned buffers from mv3d_host_alloc are then
 * DMA'd directly, pageable ones go through the driver's staging path). */
int mv3d_encode_batch_host(mv3d_ctx *ctx, const float *points_host, const int *offsets_host, int batch,
                           float *top_host, float *front_host,
                           const mv3d_top_params *tp, const mv3d_front_params *fp,
                           int semantics, int strategy);

/* ------------------------------------------------------------------------------------------------
 * Python-twin mode: the reference's PYTHON encoders, which is what src/preprocess_for_mv3d.py:30-32 and the
 * training loaders (src/utils/batch_loading.py:659-672) actually run. They differ from the C routine above
 *  and are reproduced as numpy computes them, on float32:
 *   lidar_to_top          src/data.py:296-367  -> top   [xn][yn][zn+2], written already flipped in X and Y
 *   lidar_to_front_fast   src/data.py:113-180  -> front [front_width][front_height][3], indexed [c][r], MEAN per pixel
 * The struct carries the module-level TOP_* constants (src/config.py:154-189) narrowed to float32, the grid sizes
 * as the reference derives them (xn = int((x_max-x_min)//x_div)+1 in double, zn = int((z_max-z_min)/z_div)) and the
 * front-view constants of src/config.py:32-42. filter_center_car: src/data.py:224-227 (didi data sets).
 * ---------------------------------------------------------------------------------------------- */
typedef struct mv3d_pytwin_params {
    float x_min, x_max, y_min, y_max, z_min, z_max;
    float x_div, y_div, z_div;
    int xn, yn, zn;
    int filter_center_car;
    float ang_res, vert_res, velodyne_height;
    int front_r_offset, front_c_offset, front_r_min, front_r_max, front_c_min, front_c_max;
    int front_width, front_height;
} mv3d_pytwin_params;

/* Device-resident sweeps, like mv3d_encode_batch_dev; top_dev or front_dev may be NULL to skip that view. */
int mv3d_pytwin_encode_batch_dev(mv3d_ctx *ctx, const float *points_dev, const int *offsets_host, int batch,
                                 float *top_dev, float *front_dev, const mv3d_pytwin_params *pp, void *stream);
/* Host buffers (numpy arrays): copy in, encode, copy out, pipelined sweep by sweep. */
int mv3d_pytwin_encode_batch_host(mv3d_ctx *ctx, const float *points_host, const int *offsets_host, int batch,
                                  float *top_host, float *front_host, const mv3d_pytwin_params *pp);
/* Test hook: out[i] = the device restatement of the C library's float atan2f (what np.arctan2 on float32 was). */
int mv3d_debug_atan2f(mv3d_ctx *ctx, const float *y_dev, const float *x_dev, float *out_dev, int n, void *stream);

/* Page-locked host memory for the buffers above (NULL on failure). */
void *mv3d_host_alloc(size_t bytes);
void mv3d_host_free(void *p);

/* Kernels launched by this context so far (bench.py's gpu_launches). */
long long mv3d_ctx_launch_count(const mv3d_ctx *ctx);

/* Bytes the host entry points of this context have moved over PCIe (host->device, device->host) since the last
 * reset: what bench.py reports as h2d/d2h bytes per step. */
int mv3d_ctx_transfer_bytes(mv3d_ctx *ctx, unsigned long long *h2d, unsigned long long *d2h, int reset);

/* Optional per-kernel timing for bench.py's roofline: when enabled, CUDA events are recorded on the launching
 * stream around every kernel. mv3d_ctx_kernel_times() waits for the recorded events and reports, per kernel
 * class k in [0, mv3d_kernel_count()), the accumulated device milliseconds and launch counts
 * (launch counts accumulate whether or not profiling is on). */
int mv3d_ctx_set_profiling(mv3d_ctx *ctx, int enable);
int mv3d_kernel_count(void);
const char *mv3d_kernel_name(int kernel);
int mv3d_ctx_kernel_times(mv3d_ctx *ctx, double *ms_out, long long *launches_out, int reset);

/* Test hook: the per-point decisions of the encoder, for one sweep resident in device memory.
 *   cell_dev[i]      output cell (X-1)*yn + (Y-1) of point i, or -1 when the gate rejects it
 *   pix_dev[i]       front-view pixel r*C + c through the production arithmetic, or -1 (fp may be NULL: skipped)
 *   pix_exact_dev[i] the same pixel through the reference's own expressions -- row bin in double, column bin in float on
 *                    the C library's atan2f, no fast path (the two must always agree) */
int mv3d_debug_point_keys(mv3d_ctx *ctx, const float *points_dev, int num, const mv3d_top_params *tp,
                          const mv3d_front_params *fp, int semantics, int *cell_dev, int *pix_dev,
                          int *pix_exact_dev, void *stream);

/* Per-frame checksum of device-resident views -- the consumer of a streamed run whose encoded tensors never leave the
 * GPU (BASELINE config 4: 4096 x 60.8 MB through a ring of output slots). sums_dev[f] = sum of the 32-bit words of
 * frame f modulo 2^64: exact and order-independent, so equal views give equal sums on any rank; compare with
 * numpy: view.view(np.uint32).astype(np.uint64).sum(). Enqueues on `stream` (one streaming read of the views). */
int mv3d_frame_checksums_dev(mv3d_ctx *ctx, const float *views_dev, int frames, long long floats_per_frame,
                             unsigned long long *sums_dev, void *stream);

/* The library's own bounds checks, for machines without compute-sanitizer: a build with -DMV3D_SELF_CHECK tests every
 * index its kernels compute into scratch buffers, shared-memory tiles and outputs, and records violations as bits
 * (csrc/selfcheck.cuh). *flags receives and clears them. Returns 1 when the checks are compiled in, 0 when not
 * (*flags = 0), -1 on error. The GPU test suite asserts *flags == 0 after every test when it runs on such a build. */
int mv3d_debug_selfcheck(mv3d_ctx *ctx, unsigned *flags);
/* Commits one deliberate violation (bit 15), so that a test can watch the checks report and clear; no-op otherwise. */
int mv3d_debug_selfcheck_trip(mv3d_ctx *ctx);

/* Test hooks (no GPU needed): the host half of the compact read-back used by the host-buffer entry points. `words` is
 * what the compact tile kernel reports for one sweep -- per touched cell {cell index, ceil(planes / 32) plane masks, the
 * values of the planes whose bit is set} -- and `table` holds {first word, touched cells} per tile (tile t = tile_cells
 * consecutive cells, its cells in ascending order). dst[cells][planes] is rebuilt by `threads` host threads: _sorted in
 * one pass of streaming stores (what every sweep but the first of a call takes), _split as zero-fill + scatter (the
 * first sweep: the fill overlaps the GPU's work). Whatever words/table hold, nothing is written outside dst. */
int mv3d_debug_expand_sorted(float *dst, size_t cells, int planes, const unsigned *words, unsigned n_words,
                             const unsigned *table, int tiles, int tile_cells, int threads);
int mv3d_debug_expand_split(float *dst, size_t cells, int planes, const unsigned *words, unsigned n_words,
                            const unsigned *table, int tiles, int tile_cells, int threads);

/* NO COUNTERPART IN THE REFERENCE -- added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * The front view [batch][R][C][F] of `batch` device-resident sweeps, exactly as mv3d_encode_batch_dev(ctx, ..., top_dev
 * = NULL, front_dev, ...) writes it (bit-identical: same arithmetic, the bird's-eye-view gate still applies), but with
 * NO allocation and no context: the per-pixel scratch ("last point in input order wins") lives in a static device
 * array of MV3D_NOALLOC_MAX_PIXELS 32-bit words (1 MiB, held by every GPU that loads the library until exit). The call
 * makes no cudaMalloc / cudaFree / allocating library call; it enqueues one memset and two kernels on `stream` of the
 * CURRENT device (the first call on a device also looks the array's address up: make that call outside graph capture).
 * Limits, rejected with a message in mv3d_last_error(): batch <= 64 and batch * R * C <= MV3D_NOALLOC_MAX_PIXELS
 * (e.g. 8 sweeps of 64 x 512); tp->flags must be 0. All calls on a GPU share the scratch: each call clears what it
 * uses first, but calls that OVERLAP on the GPU (different streams without ordering) race -- serialise them.
 * offsets_host as for mv3d_encode_batch_dev. Returns 0, or -1 with mv3d_last_error(). */
#define MV3D_NOALLOC_MAX_PIXELS 262144
int mv3d_front_view_noalloc(const float *points_dev, const int *offsets_host, int batch, float *front_dev,
                            const mv3d_top_params *tp, const mv3d_front_params *fp, int semantics, void *stream);

/* Diagnostic (no GPU needed): GB/s at which `threads` threads of the library's host pool zero-fill `bytes` of memory
 * with streaming stores (mean over `reps` passes) -- the ceiling of any implementation that must write the caller's dense
 * host tensor through the CPU; bench.py prints it beside the end-to-end figure. pin_sharers > 0 pins the threads to
 * share `pin_share_index` of `pin_sharers` of this process's CPUs, as the host entry points do under torchrun.
 * Returns < 0 on failure. */
double mv3d_debug_host_fill_gbs(size_t bytes, int threads, int reps, int pin_share_index, int pin_sharers);
/* The same on a buffer of the caller's (it is overwritten with zeros): the rate at which THAT memory -- e.g. the
 * page-locked tensor an end-to-end run just received -- takes streaming stores from `threads` threads. */
double mv3d_debug_host_fill_buffer_gbs(float *dst, size_t bytes, int threads, int reps, int pin_share_index, int pin_sharers);

/* The strategy a request resolves to for a parameter set (tile strategies fall back to _ATOMIC when a tile
 * cannot fit shared memory). AUTO is refined per call: _ATOMIC for device calls of up to four sweeps or when the points
 * weigh more than 7 % of the output bytes, _TILE otherwise (measured crossovers); it never picks
 * _PATCH or _SORT. */
int mv3d_resolve_strategy(const mv3d_ctx *ctx, const mv3d_top_params *tp, int strategy);

/* Library / build identification, e.g. "mv3d_encode 0.1 sm_100a". */
const char *mv3d_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MV3D_ENCODE_H */
