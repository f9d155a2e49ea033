/*
 * mv3d_anchors.h -- C ABI of the empty-anchor filter, the first consumer of the bird's-eye view
 * (SURVEY.md section 8(f) row 2). Same library as mv3d_encode.h (libmv3d_encode.so), same context.
 *
 * Replaces src/net/utility/remove_empty_box.py:25-47 `remove_empty_anchor(view, anchors, limit)` and the pycuda
 * kernel behind it (src/net/utility/remove_empty_box_kernel.cu:12-42), called once per frame in training and
 * prediction (src/mv3d.py:280,343,1139,1460). The reference uploads the whole view (800x600x27 float32, 52 MB) for
 * every call; here the view is the tensor the encoder just left in device memory.
 *
 * Semantics (all reproduced, the data race excepted):
 *   anchors[n][4] int32 = (y1, x1, y2, x2); x indexes view dimension 0, y dimension 1 (kernel :16-20).
 *   Each coordinate is clamped to [0, dim-1] (:22-29); the box is rows [x1, x2) x columns [y1, y2), and only its
 *   first MV3D_ANCHOR_MAX_LINES rows count (the launch grid is 50 rows high, remove_empty_box.py:43).
 *   Per anchor and channel: float32 row sums, left to right from 0 (`integration`, :3-10), added row after row;
 *   then numpy's float32 row sum over the channels; the anchor is kept when that total > limit.
 *   The reference adds the rows with an unsynchronised `+=` from up to 50 thread blocks (:41), so its own totals vary
 *   from run to run; this library implements the ascending-row order.
 * Output: the kept anchor indices in ascending order, like np.where(...)[0], and their number.
 */
#ifndef MV3D_ANCHORS_H
#define MV3D_ANCHORS_H

#include "mv3d_encode.h"

#ifdef __cplusplus
extern "C" {
#endif

#define MV3D_ANCHOR_MAX_LINES 50

/* How the decision is computed.
 *   _AUTO   limit == 0 and a view without negative values, -0 or NaN (checked on the device while it is read, never
 *           assumed): total > 0 <=> the box holds a positive value, so a one-bit-per-cell occupancy map decides --
 *           one streaming read of the view. Any other input takes the exact sums, selected on the device.
 *   _EXACT  always the float32 sums in the reference's order. */
enum { MV3D_ANCHORS_AUTO = 0, MV3D_ANCHORS_EXACT = 1 };

/* Upload the anchor set (src/net/rpn_target_op.py:86-116 `make_anchors`, fixed for a network) once; it stays
 * resident in the context. anchors_host: int32 [n][4]. */
int mv3d_ctx_set_anchors(mv3d_ctx *ctx, const int *anchors_host, int n);

/* Filter the resident anchors against `frames` device-resident views, views_dev[frames][xn][yn][planes] float32
 * (exactly what mv3d_encode_batch_dev / mv3d_pytwin_encode_batch_dev wrote), on `stream`, asynchronously.
 *   keep_dev   int32 [frames][n]: the first count_dev[f] entries of row f are the kept anchor indices, ascending
 *   count_dev  int32 [frames]
 *   totals_dev float32 [frames][n] or NULL: the per-anchor totals (forces the exact sums) */
int mv3d_remove_empty_anchor_dev(mv3d_ctx *ctx, const float *views_dev, int frames, int xn, int yn, int planes,
                                 float limit, int mode, int *keep_dev, int *count_dev, float *totals_dev,
                                 void *stream);

/* Encode and filter in one call: mv3d_encode_batch_dev (resp. mv3d_pytwin_encode_batch_dev) followed by
 * mv3d_remove_empty_anchor_dev on the views it wrote -- same arguments, same results -- except that the kernel that
 * builds the view also reports which cells it touched, so the filter does not read the 52 MB view again. Applies
 * when limit == 0, mode is _AUTO and the tile strategy is in use; otherwise the two steps simply run back to back.
 * top_dev is required. */
int mv3d_encode_filter_batch_dev(mv3d_ctx *ctx, const float *points_dev, const int *offsets_host, int batch,
                                 float *top_dev, float *front_dev, const mv3d_top_params *tp,
                                 const mv3d_front_params *fp, int semantics, int strategy, float limit, int mode,
                                 int *keep_dev, int *count_dev, void *stream);
int mv3d_pytwin_encode_filter_batch_dev(mv3d_ctx *ctx, const float *points_dev, const int *offsets_host, int batch,
                                        float *top_dev, float *front_dev, const mv3d_pytwin_params *pp, float limit,
                                        int mode, int *keep_dev, int *count_dev, void *stream);

/* The reference's call, with host buffers: uploads the view and the anchors, filters, returns after keep_host
 * (int32, room for n) and *count_host are complete. */
int mv3d_remove_empty_anchor_host(mv3d_ctx *ctx, const float *view_host, int xn, int yn, int planes,
                                  const int *anchors_host, int n, float limit, int mode, int *keep_host,
                                  int *count_host);

#ifdef __cplusplus
}
#endif
#endif /* MV3D_ANCHORS_H */
