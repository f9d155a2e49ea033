// mv3d_abi.cu -- the C ABI of libmv3d_encode.so (include/mv3d_encode.h): context, workspace, the batched
// device entry point, the pipelined host entry point, and the two reference symbols
// createTopMaps / createTopAndFrontMaps (ver3/LidarTopPreprocess.c:30, ver3/LidarTopAndFrontPreprocess.c:30).
//
// There is no CPU implementation in this library: every path ends in the sm_100a kernels of
// encode_kernels.cu, and every failure is reported through mv3d_last_error().
#include "ctx.h"
#include "selfcheck.cuh"
#include "../../include/mv3d_anchors.h"

#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <chrono>
#include <cstring>
#include <algorithm>
#include <memory>
#include <mutex>
#include <vector>

using namespace mv3d;

// ------------------------------------------------------------------------------------------------
// error reporting
// ------------------------------------------------------------------------------------------------
static thread_local char g_error[512] = "";

int mv3d_fail(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
    return -1;
}

extern "C" const char *mv3d_last_error(void) { return g_error; }
extern "C" const char *mv3d_version(void) { return "mv3d_encode 0.1 sm_100a"; }

static void build_density_lut_py(float *v)
{
    // src/data.py:354: min(1, np.log(count+1)/math.log(32)) -- all double, then stored into a float32 array
    for (int c = 0; c < kDensityLutSize; c++) {
        volatile double arg = (double)c + 1.0, base = 32.0;
        const double val = log(arg) / log(base);
        v[c] = (float)(val < 1.0 ? val : 1.0);
    }
}

static void build_density_lut(float *v)
{
    // Same type mix as the reference (LidarTopPreprocess.c:167): logf on a float argument, double log(32),
    // quotient in double, narrowed to float, clamped to 1 -- evaluated by this host's libm at run time.
    for (int c = 0; c < kDensityLutSize; c++) {
        volatile float arg = (float)c + 1.0f;
        volatile double base = 32.0;
        const float val = (float)((double)logf(arg) / log(base));
        v[c] = (val < 1.0f) ? val : 1.0f;
    }
}

extern "C" int mv3d_ctx_create(int device, mv3d_ctx **out)
{
    if (!out) return fail("mv3d_ctx_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail("mv3d_ctx_create: no CUDA device (%s); this library has no CPU fallback",
                    e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    if (device < 0 || device >= count) return fail("mv3d_ctx_create: device %d out of range [0,%d)", device, count);
    CUDA_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10)
        return fail("mv3d_ctx_create: device %d is sm_%d%d; this library is built for sm_100a only", device,
                    prop.major, prop.minor);
    mv3d_ctx *ctx = new (std::nothrow) mv3d_ctx();
    if (!ctx) return fail("mv3d_ctx_create: out of host memory");
    ctx->device = device;
    ctx->sm_count = prop.multiProcessorCount;
    ctx->smem_optin = prop.sharedMemPerBlockOptin;
    const char *tb = getenv("MV3D_TILE_BYTES");
    ctx->tile_bytes_target = tb ? (size_t)strtoull(tb, nullptr, 10) : (size_t)(36 * 1024);
    ctx->smem_per_sm = prop.sharedMemPerMultiprocessor;
    const char *fc = getenv("MV3D_FILL_CHUNK");
    ctx->fill_chunk = fc ? atoi(fc) : 8;
    if (ctx->fill_chunk < 1) ctx->fill_chunk = 1;
    const char *ht = getenv("MV3D_HOST_THREADS");
    ctx->host_threads = ht ? atoi(ht) : 0;
    if (const char *hs = getenv("MV3D_HOST_SPLIT_FIRST")) ctx->host_split_first = atoi(hs);
    const char *hr = getenv("MV3D_HOST_READBACK");
    ctx->host_compact = (hr && strcmp(hr, "dense") == 0) ? 0 : 1;
    if (const char *ab = getenv("MV3D_AUTO_ATOMIC_BATCH")) ctx->auto_atomic_batch = atoi(ab);
    if (const char *ar = getenv("MV3D_AUTO_ATOMIC_RATIO")) ctx->auto_atomic_ratio = atof(ar);
    int prio_lo = 0, prio_hi = 0;
    cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi);  // the zero-fill is the long pole of PATCH: never starve it
    bool ok = cudaStreamCreateWithPriority(&ctx->side_stream, cudaStreamNonBlocking, prio_hi) == cudaSuccess &&
              cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming) == cudaSuccess &&
              cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; ok && i < kMaxBatchPerLaunch; i++)
        ok = cudaEventCreateWithFlags(&ctx->ev_ready[i], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) {
        fail("mv3d_ctx_create: cannot create the pipeline stream/events: %s", cudaGetErrorString(cudaGetLastError()));
        delete ctx;
        return -1;
    }
    float lut[kDensityLutSize];
    build_density_lut(lut);
    if (cudaMalloc(&ctx->lut, sizeof(lut)) != cudaSuccess ||
        cudaMemcpy(ctx->lut, lut, sizeof(lut), cudaMemcpyHostToDevice) != cudaSuccess) {
        fail("mv3d_ctx_create: cannot allocate the density table: %s", cudaGetErrorString(cudaGetLastError()));
        delete ctx;
        return -1;
    }
    build_density_lut_py(lut);
    if (cudaMalloc(&ctx->lut_py, sizeof(lut)) != cudaSuccess ||
        cudaMemcpy(ctx->lut_py, lut, sizeof(lut), cudaMemcpyHostToDevice) != cudaSuccess) {
        fail("mv3d_ctx_create: cannot allocate the density table: %s", cudaGetErrorString(cudaGetLastError()));
        mv3d_ctx_destroy(ctx);
        return -1;
    }
    *out = ctx;
    return 0;
}

extern "C" void mv3d_ctx_destroy(mv3d_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    release(ctx->ws);
    {
        AnchorState &a = ctx->anchors;
        release(a.boxes); release(a.occupancy); release(a.flags); release(a.keep); release(a.rows);
        release(a.view); release(a.idx); release(a.count);
    }
    for (auto &s : ctx->slots) {
        release(s.ws);
        release(s.points); release(s.top); release(s.front);
        release(s.compact);
        release(s.ctable);
        if (s.stage.ptr) cudaFreeHost(s.stage.ptr);
        if (s.front_stage.ptr) cudaFreeHost(s.front_stage.ptr);
        if (s.toff_stage.ptr) cudaFreeHost(s.toff_stage.ptr);
        if (s.done) cudaEventDestroy(s.done);
        if (s.stream) cudaStreamDestroy(s.stream);
    }
    if (ctx->lut) cudaFree(ctx->lut);
    if (ctx->lut_py) cudaFree(ctx->lut_py);
    if (ctx->side_stream) cudaStreamDestroy(ctx->side_stream);
    if (ctx->ev_order) cudaEventDestroy(ctx->ev_order);
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
    for (auto e : ctx->ev_ready) if (e) cudaEventDestroy(e);
    for (auto &t : ctx->timed) { cudaEventDestroy(t.start); cudaEventDestroy(t.stop); }
    for (auto e : ctx->event_pool) cudaEventDestroy(e);
    delete ctx;
}

extern "C" long long mv3d_ctx_launch_count(const mv3d_ctx *ctx) { return ctx ? ctx->launches : 0; }

extern "C" int mv3d_ctx_transfer_bytes(mv3d_ctx *ctx, unsigned long long *h2d, unsigned long long *d2h, int reset)
{
    if (!ctx) return fail("mv3d_ctx_transfer_bytes: ctx is NULL");
    std::lock_guard<std::mutex> lock(ctx->mu);
    if (h2d) *h2d = ctx->h2d_bytes;
    if (d2h) *d2h = ctx->d2h_bytes;
    if (reset) ctx->h2d_bytes = ctx->d2h_bytes = 0;
    return 0;
}

extern "C" int mv3d_ctx_set_profiling(mv3d_ctx *ctx, int enable)
{
    if (!ctx) return fail("mv3d_ctx_set_profiling: ctx is NULL");
    std::lock_guard<std::mutex> lock(ctx->mu);
    ctx->profiling = enable != 0;
    return 0;
}

extern "C" int mv3d_kernel_count(void) { return K_COUNT; }
extern "C" const char *mv3d_kernel_name(int kernel) { return (kernel >= 0 && kernel < K_COUNT) ? kKernelNames[kernel] : ""; }

extern "C" int mv3d_ctx_kernel_times(mv3d_ctx *ctx, double *ms_out, long long *launches_out, int reset)
{
    if (!ctx) return fail("mv3d_ctx_kernel_times: ctx is NULL");
    std::lock_guard<std::mutex> lock(ctx->mu);
    CUDA_TRY(cudaSetDevice(ctx->device));
    for (auto &t : ctx->timed) {
        CUDA_TRY(cudaEventSynchronize(t.stop));
        float ms = 0.0f;
        CUDA_TRY(cudaEventElapsedTime(&ms, t.start, t.stop));
        ctx->kernel_ms[t.kernel] += ms;
        ctx->event_pool.push_back(t.start);
        ctx->event_pool.push_back(t.stop);
    }
    ctx->timed.clear();
    for (int k = 0; k < K_COUNT; k++) {
        if (ms_out) ms_out[k] = ctx->kernel_ms[k];
        if (launches_out) launches_out[k] = ctx->kernel_launches[k];
        if (reset) { ctx->kernel_ms[k] = 0; ctx->kernel_launches[k] = 0; }
    }
    return 0;
}

extern "C" void *mv3d_host_alloc(size_t bytes)
{
    void *p = nullptr;
    if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) {
        fail("mv3d_host_alloc(%zu): %s", bytes, cudaGetErrorString(cudaGetLastError()));
        return nullptr;
    }
    return p;
}

extern "C" void mv3d_host_free(void *p)
{
    if (p) cudaFreeHost(p);
}

// ------------------------------------------------------------------------------------------------
// parameter validation and kernel parameter block
// ------------------------------------------------------------------------------------------------
static bool integral(float v) { return std::isfinite(v) && v == std::floor(v); }

// Tile size for a parameter set: as many whole cells as fit the target, a multiple of 4 cells so every
// tile is a multiple of 16 bytes (bulk-store granularity) whatever the plane count.
static int pick_tile_cells(size_t bytes_target, int planes, long long cells)
{
    long long t = (long long)(bytes_target / ((size_t)planes * 4));
    t &= ~3LL;
    if (t < 4) t = 4;
    if (t > 65532) t = 65532;
    const long long cells4 = (cells + 3) & ~3LL;
    if (t > cells4) t = cells4;
    return (int)t;
}

static bool tile_feasible(const mv3d_ctx *ctx, const mv3d_top_params *tp)
{
    if (!tp || tp->zn > 65535) return false;
    const int planes = tp->zn + 2;
    const int t = pick_tile_cells(ctx->tile_bytes_target, planes, (long long)tp->xn * tp->yn);
    const size_t per = (((size_t)t * planes * 4 + 15) & ~(size_t)15) + (size_t)t * 12 + 64;
    return per <= ctx->smem_optin;
}

extern "C" int mv3d_resolve_strategy(const mv3d_ctx *ctx, const mv3d_top_params *tp, int strategy)
{
    if (strategy == MV3D_STRATEGY_ATOMIC) return MV3D_STRATEGY_ATOMIC;
    if (!ctx || !tp) return MV3D_STRATEGY_ATOMIC;
    if (strategy == MV3D_STRATEGY_SORT) return MV3D_STRATEGY_SORT;  // checked per call: the keys must fit 32 bits
    if (!tile_feasible(ctx, tp)) return MV3D_STRATEGY_ATOMIC;
    if (strategy == MV3D_STRATEGY_PATCH) return MV3D_STRATEGY_PATCH;
    return MV3D_STRATEGY_TILE;  // AUTO is refined per call by the batch size (see resolve_for_batch)
}

// AUTO, refined per call. Measured on a B200 (tools/configs_bench.py --crossover, profiles/crossover_r1j.jsonl):
//  * the tile strategy wins when the dense write dominates (64 KITTI sweeps: 78 k vs 62 k sweeps/s);
//  * up to four sweeps per call the three small launches of the atomic strategy finish first (one sweep: 53 vs 73 us);
//  * once the points weigh more than ~7 % of the output bytes (the 2 M-point cloud of config 5: 36 %) the per-tile
//    reduction chain of the hottest tiles is the long pole and the L2 atomics win (0.64 vs 0.98 ms per 8 clouds).
// points_total < 0: the caller needs the tile strategy's by-products (compact read-back, occupancy report).
static int resolve_for_batch(const mv3d_ctx *ctx, const mv3d_top_params *tp, int strategy, int batch,
                             long long points_total = -1)
{
    const int r = mv3d_resolve_strategy(ctx, tp, strategy);
    if (strategy != MV3D_STRATEGY_AUTO || r != MV3D_STRATEGY_TILE) return r;
    if (points_total >= 0 && batch > 0) {
        const double out_bytes = (double)batch * tp->xn * tp->yn * (tp->zn + 2) * 4.0;
        if (batch <= ctx->auto_atomic_batch || 16.0 * (double)points_total > ctx->auto_atomic_ratio * out_bytes)
            return MV3D_STRATEGY_ATOMIC;
    }
    return r;
}

// 1/div for the fast path of quantise() (view_math.cuh); NaN sends every point through the IEEE division.
static float fast_reciprocal(float div)
{
    const float inv = 1.0f / div;
    return (div > 0.0f && std::isfinite(inv) && inv >= 1.17549435e-38f) ? inv : std::nanf("");
}

static int make_params(const mv3d_ctx *ctx, const mv3d_top_params *tp, const mv3d_front_params *fp, bool want_top,
                       bool want_front, int semantics, ViewParams *out)
{
    if (!tp) return fail("top parameters are required (the front view is gated by the bird's-eye-view box)");
    if (semantics != MV3D_SEM_TOP && semantics != MV3D_SEM_TOPFRONT) return fail("unknown semantics %d", semantics);
    ViewParams p;
    memset(&p, 0, sizeof(p));
    if (tp->xn < 0 || tp->yn < 0 || tp->zn < 0) return fail("negative grid size %d x %d x %d", tp->xn, tp->yn, tp->zn);
    const long long cells = (long long)tp->xn * tp->yn;
    if (cells >= (1LL << 31) - 8) return fail("grid %d x %d has too many cells", tp->xn, tp->yn);
    p.x_min = tp->x_min; p.y_min = tp->y_min; p.z_min = tp->z_min;
    p.x_div = tp->x_div; p.y_div = tp->y_div; p.z_div = tp->z_div;
    p.x_inv = fast_reciprocal(tp->x_div); p.y_inv = fast_reciprocal(tp->y_div); p.z_inv = fast_reciprocal(tp->z_div);
    p.xn = tp->xn; p.yn = tp->yn; p.zn = tp->zn; p.planes = tp->zn + 2;
    p.quant_floor = (semantics == MV3D_SEM_TOPFRONT) ? 1 : 0;
    if (tp->flags & ~(MV3D_TOP_FLIP_XY | MV3D_TOP_FILTER_CENTER_CAR)) return fail("unknown top flags 0x%x", tp->flags);
    p.flip_xy = (tp->flags & MV3D_TOP_FLIP_XY) ? 1 : 0;
    p.center_car = (tp->flags & MV3D_TOP_FILTER_CENTER_CAR) ? 1 : 0;
    p.want_top = want_top ? 1 : 0;
    p.want_front = want_front ? 1 : 0;
    p.top_frame_floats = cells * p.planes;
    p.tile_cells = pick_tile_cells(ctx->tile_bytes_target, p.planes, cells);
    p.tiles_per_frame = (cells > 0 && p.tile_cells > 0) ? (int)((cells + p.tile_cells - 1) / p.tile_cells) : 0;
    p.tile_magic = tile_magic_for(p.tile_cells);
    p.z_span = tp->z_max - tp->z_min;  // float subtraction, LidarTopAndFrontPreprocess.c:185
    if (want_front) {
        if (!fp) return fail("front parameters are required when a front view is requested");
        if (!integral(fp->c_size) || !integral(fp->r_size) || !integral(fp->f_size) || fp->c_size < 1 ||
            fp->r_size < 1 || fp->f_size < 3)
            return fail("front view sizes must be integral with C,R >= 1 and F >= 3 (got C=%g R=%g F=%g)",
                        fp->c_size, fp->r_size, fp->f_size);
        const double len = (double)fp->c_size * fp->r_size * fp->f_size;
        if (len >= 16777216.0) return fail("front view of %g floats exceeds the reference's exact float index range", len);
        p.rad_to_deg = (float)(180 / 3.141596);  // sic, LidarTopAndFrontPreprocess.c:36
        p.phi_div = fp->phi_div; p.theta_div = fp->theta_div;
        p.phi_scale = p.rad_to_deg / fp->phi_div; p.theta_scale = p.rad_to_deg / fp->theta_div;
        p.phi_eps = std::fabs(p.phi_scale) * 2.5e-6f; p.theta_eps = std::fabs(p.theta_scale) * 2.5e-6f;
        p.r_size_f = fp->r_size; p.c_size_f = fp->c_size;
        p.rn = (int)fp->r_size; p.cn = (int)fp->c_size; p.fn = (int)fp->f_size;
        p.center_c = (int)(fp->c_size / 2);  // :37
        p.center_r = (int)(fp->r_size / 2);  // :38
        p.front_pixels = p.rn * p.cn;
        p.front_frame_floats = (long long)p.front_pixels * p.fn;
    }
    *out = p;
    return 0;
}

// ------------------------------------------------------------------------------------------------
// enqueue one chunk (<= kMaxBatchPerLaunch sweeps) on a stream
// ------------------------------------------------------------------------------------------------
// Resident CTAs per SM of the tile kernels for this parameter set. The shared-memory opt-in and the occupancy query
// cost tens of microseconds of host time, so they run once per (semantics, shared-memory size) and context, not per
// call; the opt-in covers every variant of the tile kernel (dense, occupancy-reporting, patch, compact).
static int tile_blocks_per_sm(mv3d_ctx *ctx, const ViewParams &p)
{
    const size_t smem = tile_smem_bytes(p);
    for (const auto &c : ctx->tile_cfg)
        if (c.smem == smem && c.py == p.py) return c.blocks_per_sm;
    int blocks = 0;
    if (tile_kernels_configure(p, &blocks) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    ctx->tile_cfg.push_back({smem, p.py, blocks});
    return blocks;
}

static int enqueue_chunk(mv3d_ctx *ctx, Workspace &ws, const float *points_dev, const int *offsets, int batch,
                         float *top_dev, float *front_dev, const ViewParams &p, int strategy, cudaStream_t s,
                         const CompactOut *compact = nullptr, const OccupancySink *sink = nullptr)
{
    FrameOffsets offs;
    memset(&offs, 0, sizeof(offs));
    const int origin = offsets[0];
    int max_points = 0;
    for (int i = 0; i <= batch; i++) offs.off[i] = offsets[i] - origin;
    for (int i = 0; i < batch; i++) {
        const int n = offs.off[i + 1] - offs.off[i];
        if (n > max_points) max_points = n;
    }
    const size_t total = (size_t)offs.off[batch];
    const float4 *pts = reinterpret_cast<const float4 *>(points_dev) + origin;
    const size_t cells = (size_t)p.xn * p.yn;

    if (p.want_front) {
        if (ensure(ws.fv_winner, (size_t)batch * p.front_pixels * 4, true, s)) return -1;
    }
    unsigned *winners = static_cast<unsigned *>(ws.fv_winner.ptr);
    const bool want_front = p.want_front && p.front_pixels > 0;

    if (p.want_top && cells > 0 && (strategy == MV3D_STRATEGY_TILE || strategy == MV3D_STRATEGY_PATCH)) {
        const bool patch = strategy == MV3D_STRATEGY_PATCH;
        if (ensure(ws.keys, total * sizeof(uint2) + 16, false, s)) return -1;
        if (ensure(ws.binned, total * sizeof(uint4) + 16, false, s)) return -1;
        const size_t hist = (size_t)batch * (p.tiles_per_frame + 1) * 4;
        if (ensure(ws.tile_count, hist, true, s)) return -1;
        if (ensure(ws.tile_off, hist, false, s)) return -1;
        const int blocks_per_sm = tile_blocks_per_sm(ctx, p);
        if (blocks_per_sm < 1) return fail("tile kernel does not fit: %zu bytes of shared memory", tile_smem_bytes(p));
        uint2 *keys = static_cast<uint2 *>(ws.keys.ptr);
        uint4 *binned = static_cast<uint4 *>(ws.binned.ptr);
        unsigned *tcount = static_cast<unsigned *>(ws.tile_count.ptr), *toff = static_cast<unsigned *>(ws.tile_off.ptr);
        const int grid_blocks = blocks_per_sm * ctx->sm_count;
        if (patch) {
            // PATCH: the copy engine zero-fills the output on the side stream (cudaMemsetAsync: no SM, 7.0-7.3 TB/s)
            // while the point passes run on the caller's stream; each chunk of sweeps is then patched (touched cells
            // only) as soon as its fill has finished. Measured comparison point of DESIGN.md 5.1, not what AUTO picks.
            const int chunk = ctx->fill_chunk;
            cudaStream_t sf = ctx->side_stream;
            CUDA_TRY(cudaEventRecord(ctx->ev_fork, s));
            CUDA_TRY(cudaStreamWaitEvent(sf, ctx->ev_fork, 0));
            for (int j0 = 0, j = 0; j0 < batch; j0 += chunk, j++) {
                const int nb = (batch - j0 < chunk) ? batch - j0 : chunk;
                float *top_j = top_dev + (size_t)j0 * (size_t)p.top_frame_floats;
                timed_launch(ctx, K_FILL_CE, sf, [&] {
                    cudaMemsetAsync(top_j, 0, (size_t)nb * (size_t)p.top_frame_floats * 4, sf);
                    return 0;  // not a kernel of this library: timed, but not counted as a launch
                });
                ctx->kernel_launches[K_FILL_CE]++;
                CUDA_TRY(cudaEventRecord(ctx->ev_ready[j], sf));
            }
            timed_launch(ctx, K_CLASSIFY, s, [&] { return launch_classify(pts, offs, p, batch, max_points, keys, tcount, winners, s); });
            timed_launch(ctx, K_TILE_SCAN, s, [&] { return launch_tile_scan(tcount, toff, p.tiles_per_frame, batch, s); });
            timed_launch(ctx, K_BIN, s, [&] {
                return launch_bin(pts, offs, p, batch, max_points, keys, toff, binned, s, want_front ? winners : nullptr, front_dev);
            });
            for (int j0 = 0, j = 0; j0 < batch; j0 += chunk, j++) {
                const int nb = (batch - j0 < chunk) ? batch - j0 : chunk;
                FrameOffsets so;
                memset(&so, 0, sizeof(so));
                for (int i = 0; i <= nb; i++) so.off[i] = offs.off[j0 + i];
                unsigned *to = toff + (size_t)j0 * (p.tiles_per_frame + 1);
                float *top_j = top_dev + (size_t)j0 * (size_t)p.top_frame_floats;
                CUDA_TRY(cudaStreamWaitEvent(s, ctx->ev_ready[j], 0));
                timed_launch(ctx, K_TILE_PATCH, s, [&] {
                    return launch_tile_patch(binned, to, so, p, ctx->lut, top_j, nb, grid_blocks, s);
                });
            }
            CUDA_TRY(cudaGetLastError());
            return 0;
        }
        // TILE: classify -> scan -> bin (+ the front view's second half, as extra CTAs of the same launch) -> tile build,
        // each launched with programmatic serialization so that its prologue overlaps the predecessor's tail.
        timed_launch(ctx, K_CLASSIFY, s, [&] { return launch_classify(pts, offs, p, batch, max_points, keys, tcount, winners, s); });
        timed_launch(ctx, K_TILE_SCAN, s, [&] { return launch_tile_scan(tcount, toff, p.tiles_per_frame, batch, s); });
        timed_launch(ctx, K_BIN, s, [&] {
            return launch_bin(pts, offs, p, batch, max_points, keys, toff, binned, s, want_front ? winners : nullptr, front_dev);
        });
        if (compact) {  // host read-back: report the touched cells only (single sweep per call)
            timed_launch(ctx, K_TILE_COMPACT, s, [&] {
                return launch_tile_compact(binned, toff, offs, p, ctx->lut, batch, grid_blocks, compact->entries,
                                           compact->cap, compact->table, s);
            });
        } else {
            timed_launch(ctx, K_TILE_BUILD, s, [&] {
                return launch_tile_build(binned, toff, offs, p, ctx->lut, top_dev, batch, grid_blocks, s, sink);
            });
        }
        CUDA_TRY(cudaGetLastError());
        return 0;
    } else if (p.want_top && cells > 0 && strategy == MV3D_STRATEGY_SORT) {
        const unsigned invalid_key = sort_invalid_key(p, batch);
        if (p.py || invalid_key == 0u)
            return fail("sort strategy: %d x %d x %d cells x %d sweeps do not fit a 32-bit key", p.xn, p.yn, p.zn, batch);
        if (ensure(ws.keys, total * 8 + 16, false, s)) return -1;
        if (ensure(ws.binned, total * 8 + 16, false, s)) return -1;
        const size_t temp_bytes = sort_temp_bytes(total, invalid_key);
        if (ensure(ws.sort_temp, temp_bytes + 16, false, s)) return -1;
        unsigned *keys_in = static_cast<unsigned *>(ws.keys.ptr), *vals_in = keys_in + total;
        unsigned *keys_out = static_cast<unsigned *>(ws.binned.ptr), *vals_out = keys_out + total;
        timed_launch(ctx, K_SORT_KEYS, s, [&] {
            return launch_sort_keys(pts, offs, p, batch, max_points, keys_in, vals_in, winners, invalid_key, s);
        });
        timed_launch(ctx, K_RADIX_SORT, s, [&] {
            return launch_sort_pairs(ws.sort_temp.ptr, temp_bytes, keys_in, keys_out, vals_in, vals_out, total, invalid_key, s);
        });
        ctx->kernel_launches[K_RADIX_SORT]++;
        timed_launch(ctx, K_FILL_ZERO, s, [&] {
            return launch_fill_zero(top_dev, (size_t)batch * (size_t)p.top_frame_floats, ctx->sm_count, s);
        });
        timed_launch(ctx, K_SORT_REDUCE, s, [&] {
            return launch_sort_reduce(pts, offs, p, ctx->lut, keys_out, vals_out, total, invalid_key, top_dev, s);
        });
    } else if (p.want_top && cells > 0) {
        if (ensure(ws.cellkey, (size_t)batch * cells * 8, false, s)) return -1;
        unsigned long long *cellkey = static_cast<unsigned long long *>(ws.cellkey.ptr);
        timed_launch(ctx, K_FILL_ZERO, s, [&] {
            return launch_fill_zero(top_dev, (size_t)batch * (size_t)p.top_frame_floats, ctx->sm_count, s) +
                   launch_fill_zero(reinterpret_cast<float *>(cellkey), (size_t)batch * cells * 2, ctx->sm_count, s);
        });
        timed_launch(ctx, K_ATOMIC_SCATTER, s, [&] {
            return launch_atomic_scatter(pts, offs, p, batch, max_points, top_dev, cellkey, winners, s);
        });
        timed_launch(ctx, K_ATOMIC_FINALIZE, s, [&] {
            return launch_atomic_finalize(pts, offs, p, ctx->lut, batch, max_points, top_dev, cellkey, s);
        });
    } else if (p.want_front) {
        // front view only: the classify pass alone elects the winners (BEV gate still applies)
        ViewParams q = p;
        q.want_top = 0;
        timed_launch(ctx, K_CLASSIFY, s, [&] { return launch_classify(pts, offs, q, batch, max_points, nullptr, nullptr, winners, s); });
    }
    if (want_front)
        timed_launch(ctx, K_FRONT_FINALIZE, s, [&] { return launch_front_finalize(pts, offs, p, batch, winners, front_dev, s); });
    CUDA_TRY(cudaGetLastError());
    return 0;
}

static int check_offsets(const int *offsets, int batch)
{
    if (batch < 0) return fail("negative batch %d", batch);
    if (!offsets) return fail("offsets is NULL");
    if (offsets[0] < 0) return fail("offsets[0] = %d is negative", offsets[0]);
    for (int i = 0; i < batch; i++) {
        if (offsets[i + 1] < offsets[i]) return fail("offsets decrease at sweep %d (%d -> %d)", i, offsets[i], offsets[i + 1]);
        if (offsets[i + 1] - offsets[i] > 1000000000) return fail("sweep %d has more than 1e9 points", i);
    }
    return 0;
}

static int encode_dev_locked(mv3d_ctx *ctx, Workspace &ws, const float *points_dev, const int *offsets, int batch,
                             float *top_dev, float *front_dev, const ViewParams &p, int strategy, cudaStream_t s,
                             const CompactOut *compact = nullptr, const OccupancySink *sink = nullptr)
{
    if (offsets[batch] > offsets[0] && (reinterpret_cast<uintptr_t>(points_dev) & 15u))
        return fail("points_dev must be 16-byte aligned");
    for (int c0 = 0; c0 < batch; c0 += kMaxBatchPerLaunch) {
        const int nb = (batch - c0 < kMaxBatchPerLaunch) ? batch - c0 : kMaxBatchPerLaunch;
        float *top_c = top_dev ? top_dev + (size_t)c0 * (size_t)p.top_frame_floats : nullptr;
        float *front_c = front_dev ? front_dev + (size_t)c0 * (size_t)p.front_frame_floats : nullptr;
        OccupancySink sink_c;
        if (sink) {
            sink_c = *sink;
            sink_c.words += (size_t)c0 * sink->words_per_frame;
            sink_c.flags += c0;
        }
        if (enqueue_chunk(ctx, ws, points_dev, offsets + c0, nb, top_c, front_c, p, strategy, s, compact,
                          sink ? &sink_c : nullptr))
            return -1;
    }
    return 0;
}

extern "C" int mv3d_encode_batch_dev(mv3d_ctx *ctx, const float *points_dev, const int *offsets_host, int batch,
                                     float *top_dev, float *front_dev, const mv3d_top_params *tp,
                                     const mv3d_front_params *fp, int semantics, int strategy, void *stream)
{
    if (!ctx) return fail("mv3d_encode_batch_dev: ctx is NULL");
    if (check_offsets(offsets_host, batch)) return -1;
    ViewParams p;
    const int strat = resolve_for_batch(ctx, tp, strategy, batch, (long long)offsets_host[batch] - offsets_host[0]);
    if (make_params(ctx, tp, fp, top_dev != nullptr, front_dev != nullptr, semantics, &p)) return -1;
    if (batch == 0 || (!top_dev && !front_dev)) return 0;
    std::lock_guard<std::mutex> lock(ctx->mu);
    CUDA_TRY(cudaSetDevice(ctx->device));
    StreamOrder order(ctx, static_cast<cudaStream_t>(stream));
    return encode_dev_locked(ctx, ctx->ws, points_dev, offsets_host, batch, top_dev, front_dev, p, strat,
                             static_cast<cudaStream_t>(stream));
}

extern "C" int mv3d_debug_point_keys(mv3d_ctx *ctx, const float *points_dev, int num, const mv3d_top_params *tp,
                                     const mv3d_front_params *fp, int semantics, int *cell_dev, int *pix_dev,
                                     int *pix_exact_dev, void *stream)
{
    if (!ctx) return fail("mv3d_debug_point_keys: ctx is NULL");
    ViewParams p;
    if (make_params(ctx, tp, fp, true, fp != nullptr, semantics, &p)) return -1;
    if (!cell_dev || (fp && (!pix_dev || !pix_exact_dev))) return fail("mv3d_debug_point_keys: output pointer is NULL");
    std::lock_guard<std::mutex> lock(ctx->mu);
    CUDA_TRY(cudaSetDevice(ctx->device));
    ctx->launches += launch_point_keys(reinterpret_cast<const float4 *>(points_dev), num, p, cell_dev, pix_dev,
                                       pix_exact_dev, static_cast<cudaStream_t>(stream));
    CUDA_TRY(cudaGetLastError());
    return 0;
}

// ------------------------------------------------------------------------------------------------
// host buffers, compact read-back: the GPU reports only the touched cells of the bird's-eye view; host threads
// rebuild the dense tensor in the caller's buffer (zero-fill + scatter of the finished values) while the GPU works
// on the next sweeps. PCIe carries ~3 MB per sweep instead of 60.8 MB; the caller's buffer may be pageable.
// ------------------------------------------------------------------------------------------------
static int py_encode_dev_locked(mv3d_ctx *ctx, Workspace &ws, const float *points_dev, const int *offsets, int batch,
                                float *top_dev, float *front_dev, const ViewParams &p, const PyFrontParams &fp,
                                cudaStream_t s, const OccupancySink *sink, const CompactOut *compact);

// Host threads of the compact read-back. Processes that share the host (one per GPU under torchrun) share its cores:
// rank r of LOCAL_WORLD_SIZE takes the r-th consecutive range of the CPUs the process may run on and pins its workers
// there (MV3D_HOST_PIN=0: unpinned), so eight ranks do not migrate over each other's caches and memory channels.
// Default thread count: the share's size, at most 16 (measured on the B200 hosts: 8 -> 16 streaming writers still
// gains 40 %, beyond that the host memory is the limit); MV3D_HOST_THREADS overrides.
static int ensure_host_pool(mv3d_ctx *ctx)
{
    if (ctx->pool) return 0;
    const char *lws = getenv("LOCAL_WORLD_SIZE"), *lr = getenv("LOCAL_RANK");
    const int sharers = lws && atoi(lws) > 0 ? atoi(lws) : 1;
    const int index = lr ? atoi(lr) : 0;
    std::vector<int> cpus = host_cpu_share(index, sharers);
    int threads = ctx->host_threads;
    if (threads <= 0) {
        size_t hw = cpus.empty() ? std::thread::hardware_concurrency() / (unsigned)sharers : cpus.size();
        if (hw < 2) hw = 2;
        threads = (int)(hw < 16 ? hw : 16);
    }
    const char *pin = getenv("MV3D_HOST_PIN");
    const bool want_pin = pin ? atoi(pin) != 0 : sharers > 1;
    ctx->pool.reset(new (std::nothrow) HostPool(threads, want_pin ? cpus : std::vector<int>()));
    if (!ctx->pool) return fail("cannot create the host thread pool");
    return 0;
}

// pyfp: non-NULL for the Python-twin mode (its own front view and encode call; everything else is shared).
static int encode_host_compact(mv3d_ctx *ctx, const float *points_host, const int *offsets_host, int batch,
                               float *top_host, float *front_host, const ViewParams &p,
                               const PyFrontParams *pyfp = nullptr)
{
    if (ensure_host_pool(ctx)) return -1;
    const size_t cells = (size_t)p.xn * p.yn;
    const size_t front_floats = pyfp ? (size_t)pyfp->pixels * 3 : (size_t)p.front_frame_floats;
    const size_t front_bytes = front_floats * 4;
    const size_t toff_words = 2 * (size_t)p.tiles_per_frame + 1;  // the tile table {first entry, entries} + the entry counter

    // A page-locked front buffer (cudaHostAlloc / cudaHostRegister) takes its copies straight from the device.
    bool front_direct = false;
    if (front_host && front_bytes) {
        cudaPointerAttributes attr;
        if (cudaPointerGetAttributes(&attr, front_host) == cudaSuccess) front_direct = attr.type == cudaMemoryTypeHost;
        else cudaGetLastError();
    }

    // The read-back stream (see tile_body, kCompact): per touched cell { index, plane masks, non-zero values }, contiguous
    // from word 0 (tile by tile in the order the tiles finished; the table says where). Its capacity covers a sweep
    // whose every point touches a cell of its own with every plane non-zero; a sweep uses a few percent of it.
    auto enqueue = [&](int i) -> int {
        HostSlot &slot = ctx->slots[i % kSlots];
        if (!slot.stream) CUDA_TRY(cudaStreamCreateWithFlags(&slot.stream, cudaStreamNonBlocking));
        if (!slot.done) CUDA_TRY(cudaEventCreateWithFlags(&slot.done, cudaEventDisableTiming));
        cudaStream_t s = slot.stream;
        const int n = offsets_host[i + 1] - offsets_host[i];
        const size_t cap = (size_t)n * ((size_t)p.planes + 1 + (size_t)((p.planes + 31) / 32));  // words
        if (cap > 0xFFFFFFF0ull) return fail("compact read-back: a sweep of %d points x %d planes does not fit a 32-bit word index", n, p.planes);
        const size_t list_bytes = cap * 4 + 16;
        if (ensure(slot.points, (size_t)n * 16 + 16, false, s)) return -1;
        if (ensure(slot.compact, list_bytes, false, s)) return -1;
        if (ensure_pinned(slot.stage, list_bytes)) return -1;
        if (ensure(slot.ctable, toff_words * 4, false, s)) return -1;
        if (ensure_pinned(slot.toff_stage, toff_words * 4)) return -1;
        if (front_host) {
            if (ensure(slot.front, front_bytes + 16, false, s)) return -1;
            if (ensure_pinned(slot.front_stage, front_bytes + 16)) return -1;
        }
        if (n > 0)
            CUDA_TRY(cudaMemcpyAsync(slot.points.ptr, points_host + (size_t)offsets_host[i] * 4, (size_t)n * 16,
                                     cudaMemcpyHostToDevice, s));
        ctx->h2d_bytes += (size_t)n * 16;
        const int local_off[2] = {0, n};
        CompactOut out;
        out.entries = static_cast<unsigned *>(slot.compact.ptr);
        out.cap = (unsigned)cap;
        out.table = static_cast<unsigned *>(slot.ctable.ptr);
        CUDA_TRY(cudaMemsetAsync(out.table + 2 * (size_t)p.tiles_per_frame, 0, 4, s));  // the sweep's entry counter
        float *front_d = front_host ? static_cast<float *>(slot.front.ptr) : nullptr;
        // `top` only has to be non-NULL here: the compact kernel never dereferences it
        const float *pts_d = static_cast<const float *>(slot.points.ptr);
        float *top_stub = reinterpret_cast<float *>(slot.compact.ptr);
        if (pyfp ? py_encode_dev_locked(ctx, slot.ws, pts_d, local_off, 1, top_stub, front_d, p, *pyfp, s, nullptr, &out)
                 : encode_dev_locked(ctx, slot.ws, pts_d, local_off, 1, top_stub, front_d, p, MV3D_STRATEGY_TILE, s, &out))
            return -1;
        CUDA_TRY(cudaMemcpyAsync(slot.toff_stage.ptr, slot.ctable.ptr, toff_words * 4, cudaMemcpyDeviceToHost, s));
        // Speculative copy sized from the previous sweep (first sweep: every slot); topped up later if short.
        size_t guess = ctx->compact_est ? (size_t)ctx->compact_est + ctx->compact_est / 4 + 4096 : cap;  // words
        if (guess > cap) guess = cap;
        slot.copied = (unsigned)guess;
        slot.cap = (unsigned)cap;
        if (guess) CUDA_TRY(cudaMemcpyAsync(slot.stage.ptr, slot.compact.ptr, guess * 4, cudaMemcpyDeviceToHost, s));
        if (front_host && front_bytes)  // page-locked caller: the DMA delivers the front view itself
            CUDA_TRY(cudaMemcpyAsync(front_direct ? static_cast<void *>(front_host + (size_t)i * front_floats) : slot.front_stage.ptr,
                                     front_d, front_bytes, cudaMemcpyDeviceToHost, s));
        ctx->d2h_bytes += toff_words * 4 + guess * 4 + (front_host ? front_bytes : 0);
        CUDA_TRY(cudaEventRecord(slot.done, s));
        return 0;
    };

    static const bool trace = getenv("MV3D_HOST_TRACE") != nullptr;
    auto now = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double t_enq = 0, t_wait = 0, t_expand = 0, t_front = 0, t0 = now();
    const int ahead = batch < kSlots ? batch : kSlots;
    for (int i = 0; i < ahead; i++)
        if (enqueue(i)) return -1;
    t_enq += now() - t0;
    for (int i = 0; i < batch; i++) {
        HostSlot &slot = ctx->slots[i % kSlots];
        // The first sweep of a call has no earlier sweep to rebuild while the GPU works (for the reference symbols,
        // one sweep per call, that is the whole call): zero-fill its view meanwhile, scatter the values afterwards.
        const bool split = i == 0 && ctx->host_split_first;
        if (split) {
            t0 = now();
            zero_fill(*ctx->pool, top_host, (size_t)p.top_frame_floats);
            t_expand += now() - t0;
        }
        t0 = now();
        CUDA_TRY(cudaEventSynchronize(slot.done));
        t_wait += now() - t0;
        const size_t cap = slot.cap;
        const unsigned *toff = static_cast<const unsigned *>(slot.toff_stage.ptr);
        const unsigned used = toff[2 * (size_t)p.tiles_per_frame];  // = words of the sweep's stream
        if (used > cap) return fail("compact read-back: %u words, capacity %zu", used, cap);
        unsigned *stage = static_cast<unsigned *>(slot.stage.ptr), *dev = static_cast<unsigned *>(slot.compact.ptr);
        if (used > slot.copied) {  // the guess was short: fetch the rest
            const size_t more = used - slot.copied;
            CUDA_TRY(cudaMemcpyAsync(stage + slot.copied, dev + slot.copied, more * 4, cudaMemcpyDeviceToHost, slot.stream));
            CUDA_TRY(cudaStreamSynchronize(slot.stream));
            ctx->d2h_bytes += more * 4;
        }
        ctx->compact_est = used;
        // While the workers rebuild sweep i, this thread delivers its front view and queues the GPU work of
        // sweep i + kSlots - 1 (the slot released by sweep i - 1: its staging was consumed one iteration ago).
        int enq_rc = 0;
        const std::function<void()> errand = [&] {
            double t1 = now();
            if (front_host && front_bytes && !front_direct)
                memcpy(front_host + (size_t)i * front_floats, slot.front_stage.ptr, front_bytes);
            t_front += now() - t1;
            t1 = now();
            if (i >= 1 && i - 1 + kSlots < batch) enq_rc = enqueue(i - 1 + kSlots);
            t_enq += now() - t1;
        };
        t0 = now();
        if (split)
            scatter_stream(*ctx->pool, top_host, cells, p.planes, stage, used, toff, p.tiles_per_frame, p.tile_cells, &errand);
        else
            expand_sorted(*ctx->pool, top_host + (size_t)i * (size_t)p.top_frame_floats, cells, p.planes, stage, used,
                          toff, p.tiles_per_frame, p.tile_cells, &errand);
        t_expand += now() - t0;
        if (enq_rc) return -1;
    }
    if (trace)
        fprintf(stderr, "[mv3d host] %d sweeps: enqueue %.2f ms, gpu wait %.2f ms, expand %.2f ms, front copy %.2f ms (%d threads)\n",
                batch, t_enq, t_wait, t_expand, t_front, ctx->pool->size());
    return 0;
}

// ------------------------------------------------------------------------------------------------
// host buffers: per-sweep pipeline over kSlots streams (copy-in, encode, copy-out overlap across sweeps)
// ------------------------------------------------------------------------------------------------
extern "C" int mv3d_encode_batch_host(mv3d_ctx *ctx, const float *points_host, const int *offsets_host, int batch,
                                      float *top_host, float *front_host, const mv3d_top_params *tp,
                                      const mv3d_front_params *fp, int semantics, int strategy)
{
    if (!ctx) return fail("mv3d_encode_batch_host: ctx is NULL");
    if (check_offsets(offsets_host, batch)) return -1;
    ViewParams p;
    const int strat = resolve_for_batch(ctx, tp, strategy, 1);  // the host pipeline encodes sweep by sweep
    if (make_params(ctx, tp, fp, top_host != nullptr, front_host != nullptr, semantics, &p)) return -1;
    if (batch == 0 || (!top_host && !front_host)) return 0;
    if (offsets_host[batch] > offsets_host[0] && !points_host) return fail("points_host is NULL");
    std::lock_guard<std::mutex> lock(ctx->mu);
    CUDA_TRY(cudaSetDevice(ctx->device));
    const size_t top_bytes = (size_t)p.top_frame_floats * 4, front_bytes = (size_t)p.front_frame_floats * 4;

    if (ctx->host_compact && top_host && top_bytes && strat == MV3D_STRATEGY_TILE) {
        int rcc = encode_host_compact(ctx, points_host, offsets_host, batch, top_host, front_host, p);
        if (rcc)  // leave no slot mid-flight behind a failure
            for (auto &slot : ctx->slots)
                if (slot.stream) cudaStreamSynchronize(slot.stream);
        return rcc;
    }

    int rc = 0;
    for (int i = 0; i < batch && rc == 0; i++) {
        HostSlot &slot = ctx->slots[i % kSlots];
        if (!slot.stream) {
            const cudaError_t se = cudaStreamCreateWithFlags(&slot.stream, cudaStreamNonBlocking);
            if (se != cudaSuccess) { slot.stream = nullptr; rc = fail("cannot create a stream: %s", cudaGetErrorString(se)); break; }
        }
        cudaStream_t s = slot.stream;
        const int n = offsets_host[i + 1] - offsets_host[i];
        // Stream order makes the slot's previous copy-out finish before its buffers are reused.
        if (ensure(slot.points, (size_t)n * 16 + 16, false, s)) { rc = -1; break; }
        if (top_host && ensure(slot.top, top_bytes + 16, false, s)) { rc = -1; break; }
        if (front_host && ensure(slot.front, front_bytes + 16, false, s)) { rc = -1; break; }
        // A failure must not return while earlier copies are still writing the caller's buffers: break, then drain.
        cudaError_t ce = cudaSuccess;
        if (n > 0)
            ce = cudaMemcpyAsync(slot.points.ptr, points_host + (size_t)offsets_host[i] * 4, (size_t)n * 16,
                                 cudaMemcpyHostToDevice, s);
        if (ce != cudaSuccess) { rc = fail("copy-in of sweep %d failed: %s", i, cudaGetErrorString(ce)); break; }
        const int local_off[2] = {0, n};
        float *top_d = top_host ? static_cast<float *>(slot.top.ptr) : nullptr;
        float *front_d = front_host ? static_cast<float *>(slot.front.ptr) : nullptr;
        if (encode_dev_locked(ctx, slot.ws, static_cast<const float *>(slot.points.ptr), local_off, 1, top_d, front_d,
                              p, strat, s)) { rc = -1; break; }
        ctx->h2d_bytes += (size_t)n * 16;
        ctx->d2h_bytes += (top_host ? top_bytes : 0) + (front_host ? front_bytes : 0);
        if (top_host && top_bytes)
            ce = cudaMemcpyAsync(top_host + (size_t)i * (size_t)p.top_frame_floats, top_d, top_bytes,
                                 cudaMemcpyDeviceToHost, s);
        if (ce == cudaSuccess && front_host && front_bytes)
            ce = cudaMemcpyAsync(front_host + (size_t)i * (size_t)p.front_frame_floats, front_d, front_bytes,
                                 cudaMemcpyDeviceToHost, s);
        if (ce != cudaSuccess) { rc = fail("copy-out of sweep %d failed: %s", i, cudaGetErrorString(ce)); break; }
    }
    for (auto &slot : ctx->slots)
        if (slot.stream) {
            cudaError_t e = cudaStreamSynchronize(slot.stream);
            if (e != cudaSuccess && rc == 0) rc = fail("pipeline failed: %s", cudaGetErrorString(e));
        }
    return rc;
}

// ------------------------------------------------------------------------------------------------
// Python-twin semantics: the reference's PYTHON encoders, src/data.py:296-367
// lidar_to_top and :113-180 lidar_to_front_fast -- what preprocess_for_mv3d.py and the training loaders run.
// ------------------------------------------------------------------------------------------------
static int make_py_params(const mv3d_ctx *ctx, const mv3d_pytwin_params *pp, bool want_top, ViewParams *out,
                          PyFrontParams *fout)
{
    if (!pp) return fail("pytwin parameters are required");
    if (pp->xn < 0 || pp->yn < 0 || pp->zn < 0) return fail("negative grid size %d x %d x %d", pp->xn, pp->yn, pp->zn);
    if (pp->zn >= (int)kNoSlice) return fail("too many height slices (%d)", pp->zn);
    const long long cells = (long long)pp->xn * pp->yn;
    if (cells >= (1LL << 31) - 8) return fail("grid %d x %d has too many cells", pp->xn, pp->yn);
    ViewParams p;
    memset(&p, 0, sizeof(p));
    p.py = 1;
    p.center_car = pp->filter_center_car ? 1 : 0;
    p.x_min = pp->x_min; p.y_min = pp->y_min; p.z_min = pp->z_min;
    p.x_max = pp->x_max; p.y_max = pp->y_max; p.z_max = pp->z_max;
    p.x_div = pp->x_div; p.y_div = pp->y_div; p.z_div = pp->z_div;
    p.xn = pp->xn; p.yn = pp->yn; p.zn = pp->zn; p.planes = pp->zn + 2;
    p.want_top = want_top ? 1 : 0;
    p.want_front = 0;  // the C routine's front view is not part of this mode
    p.top_frame_floats = cells * p.planes;
    p.tile_cells = pick_tile_cells(ctx->tile_bytes_target, p.planes, cells);
    p.tiles_per_frame = (cells > 0 && p.tile_cells > 0) ? (int)((cells + p.tile_cells - 1) / p.tile_cells) : 0;
    p.tile_magic = tile_magic_for(p.tile_cells);
    PyFrontParams f;
    memset(&f, 0, sizeof(f));
    f.ang_res = pp->ang_res; f.vert_res = pp->vert_res; f.velodyne_height = pp->velodyne_height;
    f.r_offset = pp->front_r_offset; f.c_offset = pp->front_c_offset;
    f.r_min = pp->front_r_min; f.r_max = pp->front_r_max; f.c_min = pp->front_c_min; f.c_max = pp->front_c_max;
    f.width = pp->front_width; f.height = pp->front_height;
    if (f.width < 0 || f.height < 0 || (long long)f.width * f.height > (1LL << 28))
        return fail("front view %d x %d is out of range", f.width, f.height);
    f.pixels = f.width * f.height;
    *out = p;
    *fout = f;
    return 0;
}

static int enqueue_py_chunk(mv3d_ctx *ctx, Workspace &ws, const float *points_dev, const int *offsets, int batch,
                            float *top_dev, float *front_dev, const ViewParams &p, const PyFrontParams &fp, cudaStream_t s,
                            const OccupancySink *sink = nullptr, const CompactOut *compact = nullptr)
{
    FrameOffsets offs;
    memset(&offs, 0, sizeof(offs));
    const int origin = offsets[0];
    int max_points = 0;
    for (int i = 0; i <= batch; i++) offs.off[i] = offsets[i] - origin;
    for (int i = 0; i < batch; i++) max_points = std::max(max_points, offs.off[i + 1] - offs.off[i]);
    const size_t total = (size_t)offs.off[batch];
    const float4 *pts = reinterpret_cast<const float4 *>(points_dev) + origin;
    if (ensure(ws.keys, total * sizeof(uint2) + 16, false, s)) return -1;
    if (ensure(ws.binned, total * sizeof(uint4) + 16, false, s)) return -1;
    uint2 *keys = static_cast<uint2 *>(ws.keys.ptr);
    uint4 *binned = static_cast<uint4 *>(ws.binned.ptr);

    if (top_dev && p.top_frame_floats > 0) {
        mv3d_top_params tp;
        memset(&tp, 0, sizeof(tp));
        tp.xn = p.xn; tp.yn = p.yn; tp.zn = p.zn;
        if (!tile_feasible(ctx, &tp)) return fail("python-twin mode needs the tile strategy: a tile of this grid does not fit shared memory");
        const size_t hist = (size_t)batch * (p.tiles_per_frame + 1) * 4;
        if (ensure(ws.tile_count, hist, true, s)) return -1;
        if (ensure(ws.tile_off, hist, false, s)) return -1;
        unsigned *tcount = static_cast<unsigned *>(ws.tile_count.ptr), *toff = static_cast<unsigned *>(ws.tile_off.ptr);
        const int blocks_per_sm = tile_blocks_per_sm(ctx, p);
        if (blocks_per_sm < 1) return fail("tile kernel does not fit: %zu bytes of shared memory", tile_smem_bytes(p));
        timed_launch(ctx, K_CLASSIFY, s, [&] { return launch_classify(pts, offs, p, batch, max_points, keys, tcount, nullptr, s); });
        timed_launch(ctx, K_TILE_SCAN, s, [&] { return launch_tile_scan(tcount, toff, p.tiles_per_frame, batch, s); });
        timed_launch(ctx, K_BIN, s, [&] { return launch_bin(pts, offs, p, batch, max_points, keys, toff, binned, s); });
        if (compact)  // host read-back: report the touched cells only (single sweep per call)
            timed_launch(ctx, K_TILE_COMPACT, s, [&] {
                return launch_tile_compact(binned, toff, offs, p, ctx->lut_py, batch, blocks_per_sm * ctx->sm_count,
                                           compact->entries, compact->cap, compact->table, s);
            });
        else
            timed_launch(ctx, K_TILE_BUILD, s, [&] {
                return launch_tile_build(binned, toff, offs, p, ctx->lut_py, top_dev, batch, blocks_per_sm * ctx->sm_count, s, sink);
            });
    }
    if (front_dev && fp.pixels > 0) {
        // list heads per pixel (self-cleaning, zero between calls) and one link per point (reuses the key scratch)
        if (ensure(ws.fv_winner, (size_t)batch * fp.pixels * 4, true, s)) return -1;
        unsigned *head = static_cast<unsigned *>(ws.fv_winner.ptr);
        unsigned *next = reinterpret_cast<unsigned *>(keys);
        timed_launch(ctx, K_PYFRONT_LINK, s, [&] { return launch_pyfront_link(pts, offs, p, fp, batch, max_points, head, next, s); });
        timed_launch(ctx, K_PYFRONT_REDUCE, s, [&] { return launch_pyfront_reduce(pts, offs, fp, batch, head, next, front_dev, s); });
    }
    CUDA_TRY(cudaGetLastError());
    return 0;
}

static int py_encode_dev_locked(mv3d_ctx *ctx, Workspace &ws, const float *points_dev, const int *offsets, int batch,
                                float *top_dev, float *front_dev, const ViewParams &p, const PyFrontParams &fp, cudaStream_t s,
                                const OccupancySink *sink = nullptr, const CompactOut *compact = nullptr)
{
    if (offsets[batch] > offsets[0] && (reinterpret_cast<uintptr_t>(points_dev) & 15u))
        return fail("points_dev must be 16-byte aligned");
    for (int c0 = 0; c0 < batch; c0 += kMaxBatchPerLaunch) {
        const int nb = std::min(batch - c0, kMaxBatchPerLaunch);
        float *top_c = top_dev ? top_dev + (size_t)c0 * (size_t)p.top_frame_floats : nullptr;
        float *front_c = front_dev ? front_dev + (size_t)c0 * (size_t)fp.pixels * 3 : nullptr;
        OccupancySink sink_c;
        if (sink) {
            sink_c = *sink;
            sink_c.words += (size_t)c0 * sink->words_per_frame;
            sink_c.flags += c0;
        }
        if (enqueue_py_chunk(ctx, ws, points_dev, offsets + c0, nb, top_c, front_c, p, fp, s, sink ? &sink_c : nullptr,
                             compact))
            return -1;
    }
    return 0;
}

extern "C" int mv3d_pytwin_encode_batch_dev(mv3d_ctx *ctx, const float *points_dev, const int *offsets_host, int batch,
                                            float *top_dev, float *front_dev, const mv3d_pytwin_params *pp, void *stream)
{
    if (!ctx) return fail("mv3d_pytwin_encode_batch_dev: ctx is NULL");
    if (check_offsets(offsets_host, batch)) return -1;
    ViewParams p;
    PyFrontParams fp;
    if (make_py_params(ctx, pp, top_dev != nullptr, &p, &fp)) return -1;
    if (batch == 0 || (!top_dev && !front_dev)) return 0;
    std::lock_guard<std::mutex> lock(ctx->mu);
    CUDA_TRY(cudaSetDevice(ctx->device));
    StreamOrder order(ctx, static_cast<cudaStream_t>(stream));
    return py_encode_dev_locked(ctx, ctx->ws, points_dev, offsets_host, batch, top_dev, front_dev, p, fp,
                                static_cast<cudaStream_t>(stream));
}

// ------------------------------------------------------------------------------------------------
// encode + empty-anchor filter in one call (include/mv3d_anchors.h): the tile kernel reports cell occupancy while it
// builds the view, so the filter never reads the view. Falls back to encode, then the stand-alone filter, whenever
// the occupancy map cannot decide (limit != 0, exact mode) or the strategy in use is not the tile strategy.
// ------------------------------------------------------------------------------------------------
// The fused encode + filter reports occupancy bits per tile: with a tile of a multiple of 32 cells every word of the
// bit map belongs to exactly one tile (plain stores, no atomics into a map the store stream keeps evicting from L2).
static void align_tiles_to_words(const mv3d_ctx *ctx, ViewParams *p)
{
    const long long cells = (long long)p->xn * p->yn;
    const int t = p->tile_cells;
    if (cells <= 0 || (t & 31) == 0) return;
    auto resident = [&](int tile_cells) {  // CTAs of the tile kernel an SM can hold with this tile size
        ViewParams q = *p;
        q.tile_cells = tile_cells;
        const size_t bytes = tile_smem_bytes(q);
        return bytes <= ctx->smem_optin ? (int)(ctx->smem_per_sm / (bytes + 1024)) : 0;
    };
    const int up = (t + 31) & ~31, down = t & ~31;
    int pick = t;
    if (up <= 65504 && resident(up) > 0 && resident(up) >= resident(t)) pick = up;  // same residency: take the larger
    else if (down >= 32) pick = down;
    p->tile_cells = pick;
    p->tile_magic = tile_magic_for(pick);
    p->tiles_per_frame = (int)((cells + pick - 1) / pick);
}

static int check_filter_args(const char *who, mv3d_ctx *ctx, int batch, float *top_dev, int mode, int *keep_dev,
                             int *count_dev)
{
    if (!top_dev) return fail("%s: top_dev is NULL (the filter works on the bird's-eye view)", who);
    if (mode != MV3D_ANCHORS_AUTO && mode != MV3D_ANCHORS_EXACT) return fail("%s: unknown mode %d", who, mode);
    if (batch > 0 && (!count_dev || (ctx->anchors.n > 0 && !keep_dev))) return fail("%s: NULL output buffer", who);
    if (batch > 65535) return fail("%s: at most 65535 frames per call (got %d)", who, batch);
    return 0;
}

extern "C" int mv3d_encode_filter_batch_dev(mv3d_ctx *ctx, const float *points_dev, const int *offsets_host, int batch,
                                            float *top_dev, float *front_dev, const mv3d_top_params *tp,
                                            const mv3d_front_params *fp, int semantics, int strategy, float limit,
                                            int mode, int *keep_dev, int *count_dev, void *stream)
{
    const char *who = "mv3d_encode_filter_batch_dev";
    if (!ctx) return fail("%s: ctx is NULL", who);
    if (check_offsets(offsets_host, batch)) return -1;
    if (check_filter_args(who, ctx, batch, top_dev, mode, keep_dev, count_dev)) return -1;
    ViewParams p;
    const int strat = resolve_for_batch(ctx, tp, strategy, batch);
    if (make_params(ctx, tp, fp, true, front_dev != nullptr, semantics, &p)) return -1;
    if (anchor_check_view(who, batch, p.xn, p.yn, p.planes)) return -1;
    if (batch == 0) return 0;
    std::lock_guard<std::mutex> lock(ctx->mu);
    CUDA_TRY(cudaSetDevice(ctx->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    StreamOrder order(ctx, s);
    AnchorPlan plan;
    OccupancySink sink;
    const bool fused = strat == MV3D_STRATEGY_TILE && ctx->anchors.n > 0;
    if (fused && !(limit != 0.0f) && mode == MV3D_ANCHORS_AUTO) align_tiles_to_words(ctx, &p);
    if (anchor_plan_locked(ctx, batch, p.xn, p.yn, p.planes, limit, mode, false, s, &plan, fused ? &sink : nullptr))
        return -1;
    const bool have = fused && sink.words != nullptr;
    if (encode_dev_locked(ctx, ctx->ws, points_dev, offsets_host, batch, top_dev, front_dev, p, strat, s, nullptr,
                          have ? &sink : nullptr))
        return -1;
    return anchor_run_locked(ctx, plan, top_dev, limit, keep_dev, count_dev, nullptr, s, have);
}

extern "C" int mv3d_pytwin_encode_filter_batch_dev(mv3d_ctx *ctx, const float *points_dev, const int *offsets_host,
                                                   int batch, float *top_dev, float *front_dev,
                                                   const mv3d_pytwin_params *pp, float limit, int mode, int *keep_dev,
                                                   int *count_dev, void *stream)
{
    const char *who = "mv3d_pytwin_encode_filter_batch_dev";
    if (!ctx) return fail("%s: ctx is NULL", who);
    if (check_offsets(offsets_host, batch)) return -1;
    if (check_filter_args(who, ctx, batch, top_dev, mode, keep_dev, count_dev)) return -1;
    ViewParams p;
    PyFrontParams fp;
    if (make_py_params(ctx, pp, true, &p, &fp)) return -1;
    if (anchor_check_view(who, batch, p.xn, p.yn, p.planes)) return -1;
    if (batch == 0) return 0;
    std::lock_guard<std::mutex> lock(ctx->mu);
    CUDA_TRY(cudaSetDevice(ctx->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    StreamOrder order(ctx, s);
    AnchorPlan plan;
    OccupancySink sink;
    const bool fused = ctx->anchors.n > 0;  // the Python-twin mode always runs the tile strategy
    if (fused && !(limit != 0.0f) && mode == MV3D_ANCHORS_AUTO) align_tiles_to_words(ctx, &p);
    if (anchor_plan_locked(ctx, batch, p.xn, p.yn, p.planes, limit, mode, false, s, &plan, fused ? &sink : nullptr))
        return -1;
    const bool have = fused && sink.words != nullptr;
    if (py_encode_dev_locked(ctx, ctx->ws, points_dev, offsets_host, batch, top_dev, front_dev, p, fp, s,
                             have ? &sink : nullptr))
        return -1;
    return anchor_run_locked(ctx, plan, top_dev, limit, keep_dev, count_dev, nullptr, s, have);
}

// Host buffers: sweep by sweep over the slot streams (H2D, encode, compact read-back; MV3D_HOST_READBACK=dense: dense D2H).
extern "C" int mv3d_pytwin_encode_batch_host(mv3d_ctx *ctx, const float *points_host, const int *offsets_host, int batch,
                                             float *top_host, float *front_host, const mv3d_pytwin_params *pp)
{
    if (!ctx) return fail("mv3d_pytwin_encode_batch_host: ctx is NULL");
    if (check_offsets(offsets_host, batch)) return -1;
    ViewParams p;
    PyFrontParams fp;
    if (make_py_params(ctx, pp, top_host != nullptr, &p, &fp)) return -1;
    if (batch == 0 || (!top_host && !front_host)) return 0;
    if (offsets_host[batch] > offsets_host[0] && !points_host) return fail("points_host is NULL");
    std::lock_guard<std::mutex> lock(ctx->mu);
    CUDA_TRY(cudaSetDevice(ctx->device));
    const size_t top_bytes = (size_t)p.top_frame_floats * 4, front_bytes = (size_t)fp.pixels * 3 * 4;
    if (ctx->host_compact && top_host && top_bytes) {  // compact read-back, as for the C routine's views
        int rcc = encode_host_compact(ctx, points_host, offsets_host, batch, top_host, front_host, p, &fp);
        if (rcc)
            for (auto &slot : ctx->slots)
                if (slot.stream) cudaStreamSynchronize(slot.stream);
        return rcc;
    }
    int rc = 0;
    for (int i = 0; i < batch && rc == 0; i++) {
        HostSlot &slot = ctx->slots[i % kSlots];
        if (!slot.stream) {
            const cudaError_t se = cudaStreamCreateWithFlags(&slot.stream, cudaStreamNonBlocking);
            if (se != cudaSuccess) { slot.stream = nullptr; rc = fail("cannot create a stream: %s", cudaGetErrorString(se)); break; }
        }
        cudaStream_t s = slot.stream;
        const int n = offsets_host[i + 1] - offsets_host[i];
        if (ensure(slot.points, (size_t)n * 16 + 16, false, s)) { rc = -1; break; }
        if (top_host && ensure(slot.top, top_bytes + 16, false, s)) { rc = -1; break; }
        if (front_host && ensure(slot.front, front_bytes + 16, false, s)) { rc = -1; break; }
        // A failure must not return while earlier copies are still writing the caller's buffers: break, then drain.
        cudaError_t ce = cudaSuccess;
        if (n > 0)
            ce = cudaMemcpyAsync(slot.points.ptr, points_host + (size_t)offsets_host[i] * 4, (size_t)n * 16,
                                 cudaMemcpyHostToDevice, s);
        if (ce != cudaSuccess) { rc = fail("copy-in of sweep %d failed: %s", i, cudaGetErrorString(ce)); break; }
        const int local_off[2] = {0, n};
        float *top_d = top_host ? static_cast<float *>(slot.top.ptr) : nullptr;
        float *front_d = front_host ? static_cast<float *>(slot.front.ptr) : nullptr;
        if (py_encode_dev_locked(ctx, slot.ws, static_cast<const float *>(slot.points.ptr), local_off, 1, top_d, front_d,
                                 p, fp, s)) { rc = -1; break; }
        ctx->h2d_bytes += (size_t)n * 16;
        ctx->d2h_bytes += (top_host ? top_bytes : 0) + (front_host ? front_bytes : 0);
        if (top_host && top_bytes)
            ce = cudaMemcpyAsync(top_host + (size_t)i * (size_t)p.top_frame_floats, top_d, top_bytes,
                                 cudaMemcpyDeviceToHost, s);
        if (ce == cudaSuccess && front_host && front_bytes)
            ce = cudaMemcpyAsync(front_host + (size_t)i * (size_t)fp.pixels * 3, front_d, front_bytes,
                                 cudaMemcpyDeviceToHost, s);
        if (ce != cudaSuccess) { rc = fail("copy-out of sweep %d failed: %s", i, cudaGetErrorString(ce)); break; }
    }
    for (auto &slot : ctx->slots)
        if (slot.stream) {
            cudaError_t e = cudaStreamSynchronize(slot.stream);
            if (e != cudaSuccess && rc == 0) rc = fail("pipeline failed: %s", cudaGetErrorString(e));
        }
    return rc;
}

// Test hook: the device restatement of the C library's atan2f that the Python front view uses (fd_atan2f).
__global__ void atan2f_probe_kernel(const float *y, const float *x, float *out, int n)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) out[i] = fd_atan2f(y[i], x[i]);
}

extern "C" int mv3d_debug_atan2f(mv3d_ctx *ctx, const float *y_dev, const float *x_dev, float *out_dev, int n, void *stream)
{
    if (!ctx) return fail("mv3d_debug_atan2f: ctx is NULL");
    if (n <= 0) return 0;
    std::lock_guard<std::mutex> lock(ctx->mu);
    CUDA_TRY(cudaSetDevice(ctx->device));
    atan2f_probe_kernel<<<1184, 256, 0, static_cast<cudaStream_t>(stream)>>>(y_dev, x_dev, out_dev, n);
    ctx->launches++;
    CUDA_TRY(cudaGetLastError());
    return 0;
}

// ------------------------------------------------------------------------------------------------
// per-frame checksum of device-resident views: the consumer of the streamed configuration (BASELINE config 4), where
// 4096 x 60.8 MB of output are never resident at once and never leave the GPU. Sum of the 32-bit words of a frame
// modulo 2^64: exact, independent of the order of summation, so equal views give equal sums on any rank.
// One streaming read of the views: 16-byte loads, four in flight per thread.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
frame_checksum_kernel(const float *__restrict__ views, long long floats_per_frame, unsigned long long *__restrict__ sums)
{
    const unsigned *w = reinterpret_cast<const unsigned *>(views) + (size_t)blockIdx.y * (size_t)floats_per_frame;
    unsigned long long acc = 0ull;
    // leading words up to 16-byte alignment, then vectors, then the tail
    const size_t mis = (size_t)((16u - ((unsigned)(reinterpret_cast<uintptr_t>(w)) & 15u)) & 15u) / 4;
    const size_t head = mis < (size_t)floats_per_frame ? mis : (size_t)floats_per_frame;
    const size_t nvec = ((size_t)floats_per_frame - head) / 4;
    const uint4 *v = reinterpret_cast<const uint4 *>(w + head);
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < nvec; i += 4 * stride) {
        const uint4 a = __ldcs(v + i), b = __ldcs(v + i + stride), c = __ldcs(v + i + 2 * stride), d = __ldcs(v + i + 3 * stride);
        acc += (unsigned long long)a.x + a.y + a.z + a.w;
        acc += (unsigned long long)b.x + b.y + b.z + b.w;
        acc += (unsigned long long)c.x + c.y + c.z + c.w;
        acc += (unsigned long long)d.x + d.y + d.z + d.w;
    }
    for (; i < nvec; i += stride) {
        const uint4 a = __ldcs(v + i);
        acc += (unsigned long long)a.x + a.y + a.z + a.w;
    }
    if (blockIdx.x == 0) {
        for (size_t k = threadIdx.x; k < head; k += blockDim.x) acc += w[k];
        for (size_t k = head + nvec * 4 + threadIdx.x; k < (size_t)floats_per_frame; k += blockDim.x) acc += w[k];
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xFFFFFFFFu, acc, o);
    if ((threadIdx.x & 31) == 0 && acc) atomicAdd(sums + blockIdx.y, acc);
}

extern "C" int mv3d_frame_checksums_dev(mv3d_ctx *ctx, const float *views_dev, int frames, long long floats_per_frame,
                                        unsigned long long *sums_dev, void *stream)
{
    if (!ctx) return fail("mv3d_frame_checksums_dev: ctx is NULL");
    if (frames < 0 || floats_per_frame < 0) return fail("mv3d_frame_checksums_dev: negative size");
    if (frames == 0) return 0;
    if (!sums_dev || (floats_per_frame > 0 && !views_dev)) return fail("mv3d_frame_checksums_dev: NULL buffer");
    std::lock_guard<std::mutex> lock(ctx->mu);
    CUDA_TRY(cudaSetDevice(ctx->device));
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    CUDA_TRY(cudaMemsetAsync(sums_dev, 0, (size_t)frames * sizeof(unsigned long long), s));
    if (floats_per_frame > 0) {
        for (int f0 = 0; f0 < frames; f0 += 32768) {
            const int nf = frames - f0 < 32768 ? frames - f0 : 32768;
            long long want = (floats_per_frame / 4 + 256 * 16 - 1) / (256 * 16);  // ~16 vectors per thread
            const long long cap = ((long long)ctx->sm_count * 8 + nf - 1) / nf;  // about eight CTAs per SM over all frames
            int bx = (int)(want < 1 ? 1 : (want > cap ? cap : want));
            if (bx < 1) bx = 1;
            frame_checksum_kernel<<<dim3(bx, nf), 256, 0, s>>>(views_dev + (size_t)f0 * (size_t)floats_per_frame,
                                                                floats_per_frame, sums_dev + f0);
            ctx->launches++;
        }
    }
    CUDA_TRY(cudaGetLastError());
    return 0;
}

// The library's own bounds checks (selfcheck.cuh): *flags receives and clears the violation bits the kernels of a
// -DMV3D_SELF_CHECK build have set since the last call (encode kernels in bits 0-15, sort arm shifted to 16-23).
// Returns 1 when the checks are compiled in, 0 when not (then *flags = 0), -1 on error.
extern "C" int mv3d_debug_selfcheck(mv3d_ctx *ctx, unsigned *flags)
{
    if (flags) *flags = 0;
    if (!MV3D_SELF_CHECK_ENABLED) return 0;  // whatever the arguments: a caller may probe with ctx = NULL
    if (!ctx || !flags) return fail("mv3d_debug_selfcheck: NULL argument");
    std::lock_guard<std::mutex> lock(ctx->mu);
    CUDA_TRY(cudaSetDevice(ctx->device));
    CUDA_TRY(cudaDeviceSynchronize());
    const unsigned a = selfcheck_encode(), b = selfcheck_sort();
    *flags = (a & 0xFFFFu) | ((b & 0xFFu) << 16) | ((a | b) & 0x80000000u);
    return 1;
}

// One deliberate violation, so that a test can see the checks report and clear (no-op in a normal build).
extern "C" int mv3d_debug_selfcheck_trip(mv3d_ctx *ctx)
{
    if (!ctx) return fail("mv3d_debug_selfcheck_trip: ctx is NULL");
    std::lock_guard<std::mutex> lock(ctx->mu);
    CUDA_TRY(cudaSetDevice(ctx->device));
    selfcheck_trip(nullptr);
    CUDA_TRY(cudaGetLastError());
    return 0;
}

// Test hooks for the host half of the compact read-back (need no GPU): `words` / `table` laid out as
// tile_compact_kernel writes them for one sweep (host_expand.h). _sorted is the single pass the host entry points run for
// every sweep but the first of a call, _split the two steps the first one takes (zero-fill while the GPU works, then a
// scatter of the finished values).
extern "C" int mv3d_debug_expand_sorted(float *dst, size_t cells, int planes, const unsigned *words, unsigned n_words,
                                        const unsigned *table, int tiles, int tile_cells, int threads)
{
    if (!dst || planes <= 0 || (n_words && !words) || !table || tiles <= 0 || tile_cells <= 0)
        return fail("mv3d_debug_expand_sorted: bad arguments");
    HostPool pool(threads);
    expand_sorted(pool, dst, cells, planes, words, n_words, table, tiles, tile_cells, nullptr);
    return 0;
}

extern "C" int mv3d_debug_expand_split(float *dst, size_t cells, int planes, const unsigned *words, unsigned n_words,
                                       const unsigned *table, int tiles, int tile_cells, int threads)
{
    if (!dst || planes <= 0 || (n_words && !words) || !table || tiles <= 0 || tile_cells <= 0)
        return fail("mv3d_debug_expand_split: bad arguments");
    HostPool pool(threads);
    zero_fill(pool, dst, cells * (size_t)planes);
    scatter_stream(pool, dst, cells, planes, words, n_words, table, tiles, tile_cells, nullptr);
    return 0;
}

// Diagnostic (no GPU needed): the rate at which `threads` host threads of this library's pool zero-fill `bytes` of
// fresh memory with streaming stores -- the ceiling of ANY implementation that has to write the caller's dense tensor
// through the CPU. bench.py reports it beside the end-to-end figure. Mean over `reps` passes, GB/s; < 0 on failure.
extern "C" double mv3d_debug_host_fill_gbs(size_t bytes, int threads, int reps, int pin_share_index, int pin_sharers)
{
    if (bytes < 4096 || threads < 1 || reps < 1) { fail("mv3d_debug_host_fill_gbs: bad arguments"); return -1.0; }
    void *raw = nullptr;
    if (posix_memalign(&raw, 4096, bytes) != 0) { fail("mv3d_debug_host_fill_gbs: out of memory"); return -1.0; }
    std::vector<int> cpus;
    if (pin_sharers > 0) cpus = host_cpu_share(pin_share_index, pin_sharers);
    HostPool pool(threads, cpus);
    zero_fill(pool, static_cast<float *>(raw), bytes / 4);  // first touch: page faults, not bandwidth
    // the MEAN over all passes, not the best one: ranks that share a host run this at the same time, and the best pass
    // of a rank is the one during which its neighbours happened to be idle
    const auto t0 = std::chrono::steady_clock::now();
    for (int r = 0; r < reps; r++) zero_fill(pool, static_cast<float *>(raw), bytes / 4);
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    const double best = sec > 0 ? (double)bytes * reps / sec / 1e9 : 0.0;
    free(raw);
    return best;
}

// The same measurement on a buffer of the caller's -- bench.py hands over the very host tensor its end-to-end leg just
// filled (page-locked, 4 KiB pages, already touched), so ceiling and achieved rate are measured on the same memory.
extern "C" double mv3d_debug_host_fill_buffer_gbs(float *dst, size_t bytes, int threads, int reps, int pin_share_index,
                                                  int pin_sharers)
{
    if (!dst || bytes < 4096 || threads < 1 || reps < 1) { fail("mv3d_debug_host_fill_buffer_gbs: bad arguments"); return -1.0; }
    std::vector<int> cpus;
    if (pin_sharers > 0) cpus = host_cpu_share(pin_share_index, pin_sharers);
    HostPool pool(threads, cpus);
    // In pieces of 64 MB -- about one view -- like the rebuild it is compared with: many small parts per thread, so a
    // slow core delays the end of a piece by a fraction of it.
    const size_t piece = (size_t)64 << 20;
    zero_fill(pool, dst, (piece < bytes ? piece : bytes) / 4);  // wake the pool up
    const auto t0 = std::chrono::steady_clock::now();
    for (int r = 0; r < reps; r++)
        for (size_t b = 0; b < bytes; b += piece)
            zero_fill(pool, dst + b / 4, ((bytes - b < piece) ? bytes - b : piece) / 4);
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return sec > 0 ? (double)bytes * reps / sec / 1e9 : 0.0;
}

// ------------------------------------------------------------------------------------------------
// the reference symbols
// ------------------------------------------------------------------------------------------------
// The reference routines return void and cannot fail; these can (no GPU, out of memory, sizes the CUDA path does not
// take). A failure zeroes the outputs, prints one line on stderr and leaves the message in mv3d_last_error() --
// and with MV3D_STRICT=1 in the environment it aborts the process instead of handing zeros to a caller that cannot see
// the difference (the reference's unmodified ctypes wrappers ignore everything but the buffers).
static void reference_symbol_failed(const char *who)
{
    fprintf(stderr, "libmv3d_encode: %s failed: %s\n", who, g_error);
    const char *strict = getenv("MV3D_STRICT");
    if (strict && atoi(strict) != 0) {
        fprintf(stderr, "libmv3d_encode: MV3D_STRICT is set, aborting\n");
        abort();
    }
}

static std::mutex g_default_mu;
static mv3d_ctx *g_default_ctx = nullptr;

// Created on first use, so a process that forks workers before encoding (multiprocessing.Pool in
// src/data.py:495) gets one CUDA context per worker, like the reference's pycuda loaders do.
static mv3d_ctx *default_ctx()
{
    std::lock_guard<std::mutex> lock(g_default_mu);
    if (!g_default_ctx) {
        const char *dev = getenv("MV3D_DEVICE");
        if (mv3d_ctx_create(dev ? atoi(dev) : 0, &g_default_ctx)) g_default_ctx = nullptr;
    }
    return g_default_ctx;
}

static void zero_host(void *p, double floats)
{
    if (p && floats > 0 && floats < 1e12) memset(p, 0, (size_t)floats * 4);
}

extern "C" void createTopMaps(const void *raw_data, int num, const void *top_data, float x_MIN, float x_MAX,
                              float y_MIN, float y_MAX, float z_MIN, float z_MAX, float x_DIVISION, float y_DIVISION,
                              float z_DIVISION, int X_SIZE, int Y_SIZE, int Z_SIZE)
{
    g_error[0] = 0;
    const mv3d_top_params tp = {x_MIN, x_MAX, y_MIN, y_MAX, z_MIN, z_MAX, x_DIVISION, y_DIVISION, z_DIVISION,
                                X_SIZE, Y_SIZE, Z_SIZE};
    const int offsets[2] = {0, num > 0 ? num : 0};
    mv3d_ctx *ctx = default_ctx();
    int rc = ctx ? mv3d_encode_batch_host(ctx, static_cast<const float *>(raw_data), offsets, 1,
                                          static_cast<float *>(const_cast<void *>(top_data)), nullptr, &tp, nullptr,
                                          MV3D_SEM_TOP, MV3D_STRATEGY_AUTO)
                 : -1;
    if (rc) {
        reference_symbol_failed("createTopMaps");
        if (X_SIZE > 0 && Y_SIZE > 0 && Z_SIZE >= 0)
            zero_host(const_cast<void *>(top_data), (double)X_SIZE * Y_SIZE * (Z_SIZE + 2));
    }
}

extern "C" void createTopAndFrontMaps(const void *raw_data, int num, const void *top_data, const void *front_data,
                                      float x_MIN, float x_MAX, float y_MIN, float y_MAX, float z_MIN, float z_MAX,
                                      float x_DIVISION, float y_DIVISION, float z_DIVISION, int X_SIZE, int Y_SIZE,
                                      int Z_SIZE, float phi_MIN, float phi_MAX, float theta_MIN, float theta_MAX,
                                      float phi_DIVISION, float theta_DIVISION, float C_SIZE, float R_SIZE, float F_SIZE)
{
    g_error[0] = 0;
    const mv3d_top_params tp = {x_MIN, x_MAX, y_MIN, y_MAX, z_MIN, z_MAX, x_DIVISION, y_DIVISION, z_DIVISION,
                                X_SIZE, Y_SIZE, Z_SIZE};
    const mv3d_front_params fp = {phi_MIN, phi_MAX, theta_MIN, theta_MAX, phi_DIVISION, theta_DIVISION,
                                  C_SIZE, R_SIZE, F_SIZE};
    const int offsets[2] = {0, num > 0 ? num : 0};
    mv3d_ctx *ctx = default_ctx();
    int rc = ctx ? mv3d_encode_batch_host(ctx, static_cast<const float *>(raw_data), offsets, 1,
                                          static_cast<float *>(const_cast<void *>(top_data)),
                                          static_cast<float *>(const_cast<void *>(front_data)), &tp, &fp,
                                          MV3D_SEM_TOPFRONT, MV3D_STRATEGY_AUTO)
                 : -1;
    if (rc) {
        reference_symbol_failed("createTopAndFrontMaps");
        if (X_SIZE > 0 && Y_SIZE > 0 && Z_SIZE >= 0)
            zero_host(const_cast<void *>(top_data), (double)X_SIZE * Y_SIZE * (Z_SIZE + 2));
        // a caller that hands over non-integral sizes allocated for their integer parts at most: never zero more
        if (C_SIZE >= 1 && R_SIZE >= 1 && F_SIZE >= 1 && std::isfinite(C_SIZE) && std::isfinite(R_SIZE) && std::isfinite(F_SIZE))
            zero_host(const_cast<void *>(front_data), std::floor((double)C_SIZE) * std::floor((double)R_SIZE) * std::floor((double)F_SIZE));
    }
}
