// anchor_abi.cu -- C ABI of the empty-anchor filter (include/mv3d_anchors.h) over the kernels of anchor_kernels.cu.
// Replaces src/net/utility/remove_empty_box.py:25-47. No CPU implementation: every path ends in the sm_100a kernels.
#include "../../include/mv3d_anchors.h"
#include "anchor_kernels.h"
#include "ctx.h"

#include <cmath>
#include <cstring>

static int check_view(const char *who, int frames, int xn, int yn, int planes)
{
    if (frames < 0) return fail("%s: frames = %d", who, frames);
    if (xn <= 0 || yn <= 0 || planes <= 0) return fail("%s: view shape %d x %d x %d", who, xn, yn, planes);
    if (planes > kAnchorMaxPlanes)
        return fail("%s: %d channels; the reference launches one thread per channel, so at most %d", who, planes,
                    kAnchorMaxPlanes);
    if ((long long)xn * yn >= (1ll << 31)) return fail("%s: %d x %d cells do not fit the cell index", who, xn, yn);
    return 0;
}

extern "C" int mv3d_ctx_set_anchors(mv3d_ctx *ctx, const int *anchors_host, int n)
{
    if (!ctx) return fail("mv3d_ctx_set_anchors: ctx is NULL");
    if (n < 0 || (n > 0 && !anchors_host)) return fail("mv3d_ctx_set_anchors: bad arguments (n = %d)", n);
    std::lock_guard<std::mutex> lock(ctx->mu);
    CUDA_TRY(cudaSetDevice(ctx->device));
    AnchorState &st = ctx->anchors;
    if (n == st.n && (n == 0 || std::memcmp(st.host_copy.data(), anchors_host, (size_t)n * sizeof(int4)) == 0))
        return 0;  // the set a network uses is fixed (src/mv3d.py:230): uploaded once
    CUDA_TRY(cudaDeviceSynchronize());  // rare: nothing may still read the old set
    st.n = 0;
    if (n > 0) {
        if (ensure(st.boxes, (size_t)n * sizeof(int4), false, nullptr)) return -1;
        CUDA_TRY(cudaMemcpy(st.boxes.ptr, anchors_host, (size_t)n * sizeof(int4), cudaMemcpyHostToDevice));
        ctx->h2d_bytes += (size_t)n * sizeof(int4);
        st.host_copy.assign(anchors_host, anchors_host + (size_t)n * 4);
    }
    st.n = n;
    return 0;
}

// Size and clear the scratch of one filter call over `frames` views (context lock held). With want_sink the
// occupancy map is cleared too and *sink tells the tile kernel where to report into it.
int anchor_plan_locked(mv3d_ctx *ctx, int frames, int xn, int yn, int planes, float limit, int mode, bool want_totals,
                       cudaStream_t s, AnchorPlan *plan, OccupancySink *sink)
{
    AnchorState &st = ctx->anchors;
    AnchorGeometry &g = plan->g;
    g.xn = xn;
    g.yn = yn;
    g.planes = planes;
    g.cells = (unsigned)xn * (unsigned)yn;
    g.words = (g.cells + 31u) / 32u;
    g.wpr = ((unsigned)yn + 31u) / 32u;
    g.n = st.n;
    g.chunks = (unsigned)((st.n + kAnchorChunk - 1) / kAnchorChunk);
    g.max_lines = MV3D_ANCHOR_MAX_LINES;
    // The occupancy map decides only "total > 0"; any other threshold, a NaN threshold or a request for the totals
    // themselves takes the sums for every frame.
    plan->exact_all = mode == MV3D_ANCHORS_EXACT || want_totals || !(limit == 0.0f);
    plan->frames = frames;
    if (frames == 0 || st.n == 0) return 0;
    // flags [frames] and chunk counts [frames][chunks] share one buffer so that one memset clears both
    const size_t flag_words = (size_t)frames, count_words = (size_t)frames * g.chunks;
    if (ensure(st.flags, (flag_words + count_words) * sizeof(unsigned), false, s)) return -1;
    if (ensure(st.keep, (size_t)frames * st.n, false, s)) return -1;
    plan->flags = static_cast<unsigned *>(st.flags.ptr);
    plan->chunk_count = plan->flags + flag_words;
    plan->keep = static_cast<unsigned char *>(st.keep.ptr);
    CUDA_TRY(cudaMemsetAsync(plan->flags, 0, (flag_words + count_words) * sizeof(unsigned), s));
    plan->occupancy = nullptr;
    if (!plan->exact_all) {
        if (ensure(st.occupancy, (size_t)frames * g.words * sizeof(unsigned), false, s)) return -1;
        plan->occupancy = static_cast<unsigned *>(st.occupancy.ptr);
        const size_t row_words = (size_t)frames * xn * g.wpr, group_words = (size_t)frames * ((xn + 7) / 8) * g.wpr;
        if (ensure(st.rows, (row_words + group_words) * sizeof(unsigned), false, s)) return -1;
        plan->rows = static_cast<unsigned *>(st.rows.ptr);
        plan->groups = plan->rows + row_words;
        if (sink) {
            CUDA_TRY(cudaMemsetAsync(plan->occupancy, 0, (size_t)frames * g.words * sizeof(unsigned), s));
            sink->words = plan->occupancy;
            sink->flags = plan->flags;
            sink->words_per_frame = g.words;
        }
    }
    return 0;
}

// Enqueue the filter on `s`. have_occupancy: the tile kernel already filled the occupancy map and the flags while
// it built these views, so the view is not read again.
int anchor_run_locked(mv3d_ctx *ctx, const AnchorPlan &plan, const float *views_dev, float limit, int *keep_dev,
                      int *count_dev, float *totals_dev, cudaStream_t s, bool have_occupancy)
{
    AnchorState &st = ctx->anchors;
    const AnchorGeometry &g = plan.g;
    const int frames = plan.frames;
    if (frames == 0) return 0;
    if (st.n == 0) {
        CUDA_TRY(cudaMemsetAsync(count_dev, 0, (size_t)frames * sizeof(int), s));
        return 0;
    }
    const int4 *boxes = static_cast<const int4 *>(st.boxes.ptr);
    if (!plan.exact_all) {
        if (!have_occupancy)
            timed_launch(ctx, K_ANCHOR_OCCUPANCY, s, [&] {
                return launch_anchor_occupancy(views_dev, g, frames, plan.occupancy, plan.flags, ctx->sm_count, s);
            });
        timed_launch(ctx, K_ANCHOR_TEST, s, [&] {
            return launch_anchor_test(boxes, g, frames, plan.occupancy, plan.flags, plan.rows, plan.groups, plan.keep,
                                      plan.chunk_count, s);
        });
    }
    // In AUTO mode this launch exits at once for every frame the occupancy map could decide (the usual case).
    timed_launch(ctx, K_ANCHOR_SUM, s, [&] {
        return launch_anchor_sum(views_dev, boxes, g, frames, limit, plan.flags, plan.exact_all, plan.keep,
                                 plan.chunk_count, totals_dev, ctx->sm_count, s);
    });
    timed_launch(ctx, K_ANCHOR_COMPACT, s, [&] {
        return launch_anchor_compact(g, frames, plan.keep, plan.chunk_count, keep_dev, count_dev, s);
    });
    CUDA_TRY(cudaGetLastError());
    return 0;
}

int anchor_check_view(const char *who, int frames, int xn, int yn, int planes) { return check_view(who, frames, xn, yn, planes); }

static int filter_locked(mv3d_ctx *ctx, const float *views_dev, int frames, int xn, int yn, int planes, float limit,
                         int mode, int *keep_dev, int *count_dev, float *totals_dev, cudaStream_t s)
{
    AnchorPlan plan;
    if (anchor_plan_locked(ctx, frames, xn, yn, planes, limit, mode, totals_dev != nullptr, s, &plan, nullptr)) return -1;
    return anchor_run_locked(ctx, plan, views_dev, limit, keep_dev, count_dev, totals_dev, s, false);
}

extern "C" int mv3d_remove_empty_anchor_dev(mv3d_ctx *ctx, const float *views_dev, int frames, int xn, int yn,
                                            int planes, float limit, int mode, int *keep_dev, int *count_dev,
                                            float *totals_dev, void *stream)
{
    if (!ctx) return fail("mv3d_remove_empty_anchor_dev: ctx is NULL");
    if (check_view("mv3d_remove_empty_anchor_dev", frames, xn, yn, planes)) return -1;
    if (mode != MV3D_ANCHORS_AUTO && mode != MV3D_ANCHORS_EXACT)
        return fail("mv3d_remove_empty_anchor_dev: unknown mode %d", mode);
    if (frames > 0 && (!views_dev || !count_dev || (ctx->anchors.n > 0 && !keep_dev)))
        return fail("mv3d_remove_empty_anchor_dev: NULL buffer");
    if (frames > 65535) return fail("mv3d_remove_empty_anchor_dev: at most 65535 frames per call (got %d)", frames);
    std::lock_guard<std::mutex> lock(ctx->mu);
    CUDA_TRY(cudaSetDevice(ctx->device));
    StreamOrder order(ctx, static_cast<cudaStream_t>(stream));
    return filter_locked(ctx, views_dev, frames, xn, yn, planes, limit, mode, keep_dev, count_dev, totals_dev,
                         static_cast<cudaStream_t>(stream));
}

extern "C" int mv3d_remove_empty_anchor_host(mv3d_ctx *ctx, const float *view_host, int xn, int yn, int planes,
                                             const int *anchors_host, int n, float limit, int mode, int *keep_host,
                                             int *count_host)
{
    if (!ctx) return fail("mv3d_remove_empty_anchor_host: ctx is NULL");
    if (check_view("mv3d_remove_empty_anchor_host", 1, xn, yn, planes)) return -1;
    if (!view_host || !count_host || n < 0 || (n > 0 && (!anchors_host || !keep_host)))
        return fail("mv3d_remove_empty_anchor_host: bad arguments");
    if (mode != MV3D_ANCHORS_AUTO && mode != MV3D_ANCHORS_EXACT)
        return fail("mv3d_remove_empty_anchor_host: unknown mode %d", mode);
    *count_host = 0;
    if (mv3d_ctx_set_anchors(ctx, anchors_host, n)) return -1;
    if (n == 0) return 0;
    std::lock_guard<std::mutex> lock(ctx->mu);
    CUDA_TRY(cudaSetDevice(ctx->device));
    AnchorState &st = ctx->anchors;
    cudaStream_t s = nullptr;  // the legacy default stream: this call returns only when its results are on the host
    StreamOrder order(ctx, s);
    const size_t view_bytes = (size_t)xn * yn * planes * sizeof(float);
    if (ensure(st.view, view_bytes, false, s)) return -1;
    if (ensure(st.idx, (size_t)n * sizeof(int), false, s)) return -1;
    if (ensure(st.count, sizeof(int), false, s)) return -1;
    CUDA_TRY(cudaMemcpyAsync(st.view.ptr, view_host, view_bytes, cudaMemcpyHostToDevice, s));
    ctx->h2d_bytes += view_bytes;
    if (filter_locked(ctx, static_cast<const float *>(st.view.ptr), 1, xn, yn, planes, limit, mode,
                      static_cast<int *>(st.idx.ptr), static_cast<int *>(st.count.ptr), nullptr, s))
        return -1;
    int kept = 0;
    CUDA_TRY(cudaMemcpyAsync(&kept, st.count.ptr, sizeof(int), cudaMemcpyDeviceToHost, s));
    CUDA_TRY(cudaStreamSynchronize(s));
    if (kept < 0 || kept > n) return fail("mv3d_remove_empty_anchor_host: device reported %d of %d anchors", kept, n);
    if (kept > 0) {
        CUDA_TRY(cudaMemcpyAsync(keep_host, st.idx.ptr, (size_t)kept * sizeof(int), cudaMemcpyDeviceToHost, s));
        CUDA_TRY(cudaStreamSynchronize(s));
    }
    ctx->d2h_bytes += sizeof(int) + (size_t)kept * sizeof(int);
    *count_host = kept;
    return 0;
}
