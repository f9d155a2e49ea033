// ctx.h -- the context behind the C ABI (internal; shared by mv3d_abi.cu and anchor_abi.cu).
#pragma once
#include "../../include/mv3d_encode.h"
#include "anchor_kernels.h"
#include "encode_kernels.h"
#include "host_expand.h"
#include "view_math.cuh"

#include <cuda_runtime.h>
#include <memory>
#include <mutex>
#include <vector>

// Records the message mv3d_last_error() returns on this thread; always returns -1.
int mv3d_fail(const char *fmt, ...);
#define fail mv3d_fail

#define CUDA_TRY(expr)                                                                              \
    do {                                                                                            \
        cudaError_t _e = (expr);                                                                    \
        if (_e != cudaSuccess)                                                                      \
            return fail("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
    } while (0)

using namespace mv3d;

// ------------------------------------------------------------------------------------------------
// context
// ------------------------------------------------------------------------------------------------
struct DeviceBuffer {
    void *ptr = nullptr;
    size_t bytes = 0;
};

// Scratch for one in-flight encode call. Zero-initialised where the kernels rely on it (tile histogram,
// front-view winners) and left zero by them, so no per-call memset is needed on the tile path.
struct Workspace {
    DeviceBuffer keys, binned, tile_count, tile_off, fv_winner, cellkey, sort_temp;
};

constexpr int kSlots = 4;  // host pipeline depth: copy-in / encode / copy-out / host rebuild of consecutive sweeps overlap

// Page-locked host staging that grows on demand.
struct PinnedBuffer {
    void *ptr = nullptr;
    size_t bytes = 0;
};

// Where the compact variant of the tile kernel reports the touched cells of ONE sweep (device memory).
struct CompactOut {
    unsigned *entries = nullptr;  // the sweep's stream of words: per touched cell {index, plane masks, non-zero values}
    unsigned cap = 0;             // its capacity in words: >= points x (planes + 1 + ceil(planes / 32))
    unsigned *table = nullptr;    // {first word, touched cells} per tile, then the sweep's word counter (zeroed before the launch)
};

// Empty-anchor filter (anchor_abi.cu): the anchor set stays resident, scratch grows on demand.
struct AnchorState {
    DeviceBuffer boxes;       // int4 [n]: (y1, x1, y2, x2) as the reference passes them
    std::vector<int> host_copy;  // what `boxes` holds: an unchanged set is not uploaded again
    int n = 0;
    DeviceBuffer occupancy;   // uint32 [frames][ceil(cells/32)]: bit = cell holds a positive value
    DeviceBuffer flags;       // uint32 [frames]: view holds a negative value, a -0 or a NaN (exact sums needed),
                              // followed by the kept-anchor counts uint32 [frames][ceil(n/1024)]
    DeviceBuffer rows;        // uint32 [frames][xn][ceil(yn/32)] row-aligned copy of the map, then its 8-row ORs
    DeviceBuffer keep;        // uint8  [frames][n]
    DeviceBuffer view, idx, count;  // staging of the host entry point
};

struct HostSlot {
    cudaStream_t stream = nullptr;
    Workspace ws;
    DeviceBuffer points, top, front;
    // compact read-back
    DeviceBuffer compact, ctable;  // the sweep's touched cells; its tile table + entry counter
    PinnedBuffer stage, front_stage, toff_stage;  // toff_stage receives the tile table
    cudaEvent_t done = nullptr;
    unsigned copied = 0;          // words already copied to `stage` by the speculative D2H
    unsigned cap = 0;             // capacity in words of the stream of the sweep in flight
};

// Kernel classes, in the order mv3d_kernel_name() reports them.
enum KernelId { K_CLASSIFY = 0, K_TILE_SCAN, K_BIN, K_TILE_BUILD, K_FRONT_FINALIZE, K_FILL_ZERO, K_ATOMIC_SCATTER,
                K_ATOMIC_FINALIZE, K_TILE_COMPACT, K_PYFRONT_LINK,
                K_PYFRONT_REDUCE, K_ANCHOR_OCCUPANCY, K_ANCHOR_TEST, K_ANCHOR_SUM, K_ANCHOR_COMPACT, K_FILL_CE,
                K_TILE_PATCH, K_SORT_KEYS, K_RADIX_SORT, K_SORT_REDUCE, K_COUNT };
static const char *const kKernelNames[K_COUNT] = {"classify_kernel", "tile_scan_kernel", "bin_kernel",
                                                  "tile_build_kernel", "front_finalize_kernel", "fill_zero_kernel",
                                                  "atomic_scatter_kernel", "atomic_finalize_kernel",
                                                  "tile_compact_kernel",
                                                  "pyfront_link_kernel", "pyfront_reduce_kernel",
                                                  "anchor_occupancy_kernel", "anchor_test_kernel",
                                                  "anchor_sum_kernel", "anchor_compact_kernel",
                                                  "ce_zero_fill", "tile_patch_kernel", "sort_keys_kernel",
                                                  "cub_radix_sort", "sort_reduce_kernel"};

struct TimedLaunch {
    cudaEvent_t start, stop;
    int kernel;
};

struct mv3d_ctx {
    bool profiling = false;
    std::vector<TimedLaunch> timed;        // pending (not yet accumulated) event pairs
    std::vector<cudaEvent_t> event_pool;   // recycled events
    double kernel_ms[K_COUNT] = {0};
    long long kernel_launches[K_COUNT] = {0};
    int device = 0;
    int sm_count = 0;
    size_t smem_optin = 0;
    float *lut = nullptr;     // density table of the C routines
    float *lut_py = nullptr;  // density table of the Python encoder (all-double log)
    AnchorState anchors;
    Workspace ws;            // used by mv3d_encode_batch_dev (one call in flight per context)
    HostSlot slots[kSlots];  // used by mv3d_encode_batch_host
    std::unique_ptr<mv3d::HostPool> pool;  // host threads of the compact read-back (created on first use)
    int host_threads = 0;
    int host_split_first = 1; // first sweep of a host call: zero-fill while the GPU encodes, scatter afterwards (MV3D_HOST_SPLIT_FIRST=0: one pass)
    int host_compact = 1;    // MV3D_HOST_READBACK=dense switches the host entry points back to a dense D2H copy
    unsigned long long h2d_bytes = 0, d2h_bytes = 0;  // moved by the host entry points since the last reset
    unsigned compact_est = 0;  // entries the previous sweep produced: sizes the speculative copy of the next one
    long long launches = 0;
    size_t tile_bytes_target = 0;
    int fill_chunk = 0;                   // PATCH: sweeps per copy-engine zero-fill
    int auto_atomic_batch = 4;            // AUTO: up to this many sweeps per device call go to the atomic strategy
    double auto_atomic_ratio = 0.07;      // AUTO: ... and so do calls whose points weigh more than this share of the output bytes
    size_t smem_per_sm = 0;
    struct TileConfig { size_t smem; int py; int blocks_per_sm; };
    std::vector<TileConfig> tile_cfg;     // shared-memory opt-in + occupancy of the tile kernels, per (size, semantics)
    cudaStream_t side_stream = nullptr;   // copy-engine zero-fill of the PATCH strategy
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_ready[kMaxBatchPerLaunch] = {nullptr};
    cudaEvent_t ev_order = nullptr;       // end of the latest device call on this context (see StreamOrder)
    cudaStream_t order_stream = nullptr;
    bool order_valid = false;
    std::mutex mu;
};

// Successive device calls on one context share its scratch buffers, so their work must not overlap on the device.
// Calls on the same stream are ordered anyway; a call on another stream first waits (on the device, not the host) for
// the end of the previous call. Construct under ctx->mu, after cudaSetDevice.
struct StreamOrder {
    mv3d_ctx *ctx;
    cudaStream_t s;
    StreamOrder(mv3d_ctx *c, cudaStream_t stream) : ctx(c), s(stream)
    {
        if (!ctx->ev_order) cudaEventCreateWithFlags(&ctx->ev_order, cudaEventDisableTiming);
        if (ctx->ev_order && ctx->order_valid && ctx->order_stream != s) cudaStreamWaitEvent(s, ctx->ev_order, 0);
    }
    ~StreamOrder()
    {
        if (!ctx->ev_order) return;
        cudaEventRecord(ctx->ev_order, s);
        ctx->order_stream = s;
        ctx->order_valid = true;
    }
};

inline int ensure(DeviceBuffer &b, size_t bytes, bool zero, cudaStream_t s)
{
    if (bytes <= b.bytes) return 0;
    // Growing is rare (first call / larger batch); make sure nothing still uses the old allocation.
    CUDA_TRY(cudaStreamSynchronize(s));
    if (b.ptr) CUDA_TRY(cudaFree(b.ptr));
    b.ptr = nullptr;
    b.bytes = 0;
    size_t want = bytes + bytes / 8;
    want = (want + 255) & ~(size_t)255;
    CUDA_TRY(cudaMalloc(&b.ptr, want));
    b.bytes = want;
    if (zero) {
        CUDA_TRY(cudaMemsetAsync(b.ptr, 0, want, s));
        CUDA_TRY(cudaStreamSynchronize(s));
    }
    return 0;
}

inline int ensure_pinned(PinnedBuffer &b, size_t bytes)
{
    if (bytes <= b.bytes) return 0;
    if (b.ptr) CUDA_TRY(cudaFreeHost(b.ptr));
    b.ptr = nullptr;
    b.bytes = 0;
    size_t want = bytes + bytes / 4;
    want = (want + 4095) & ~(size_t)4095;
    CUDA_TRY(cudaHostAlloc(&b.ptr, want, cudaHostAllocDefault));
    b.bytes = want;
    return 0;
}

inline void release(DeviceBuffer &b)
{
    if (b.ptr) cudaFree(b.ptr);
    b.ptr = nullptr;
    b.bytes = 0;
}

inline void release(Workspace &w)
{
    release(w.keys); release(w.binned); release(w.tile_count); release(w.tile_off);
    release(w.fv_winner); release(w.cellkey); release(w.sort_temp);
}

// Optional per-kernel timing: CUDA events recorded on the launching stream around each kernel.
inline cudaEvent_t take_event(mv3d_ctx *ctx)
{
    cudaEvent_t e = nullptr;
    if (!ctx->event_pool.empty()) {
        e = ctx->event_pool.back();
        ctx->event_pool.pop_back();
    } else if (cudaEventCreate(&e) != cudaSuccess) {
        e = nullptr;
    }
    return e;
}

template <typename F>
inline void timed_launch(mv3d_ctx *ctx, int kernel, cudaStream_t s, F &&launch)
{
    TimedLaunch t{nullptr, nullptr, kernel};
    if (ctx->profiling) {
        t.start = take_event(ctx);
        t.stop = take_event(ctx);
        if (t.start && t.stop) cudaEventRecord(t.start, s);
    }
    const int n = launch();
    ctx->launches += n;
    ctx->kernel_launches[kernel] += n;
    if (ctx->profiling && t.start && t.stop) {
        cudaEventRecord(t.stop, s);
        ctx->timed.push_back(t);
    }
}

// Empty-anchor filter internals shared with the fused encode+filter entry points (anchor_abi.cu).
struct AnchorPlan {
    AnchorGeometry g;
    int frames = 0;
    bool exact_all = false;
    unsigned *flags = nullptr, *chunk_count = nullptr, *occupancy = nullptr, *rows = nullptr, *groups = nullptr;
    unsigned char *keep = nullptr;
};
int anchor_check_view(const char *who, int frames, int xn, int yn, int planes);
int anchor_plan_locked(mv3d_ctx *ctx, int frames, int xn, int yn, int planes, float limit, int mode, bool want_totals,
                       cudaStream_t s, AnchorPlan *plan, OccupancySink *sink);
int anchor_run_locked(mv3d_ctx *ctx, const AnchorPlan &plan, const float *views_dev, float limit, int *keep_dev,
                      int *count_dev, float *totals_dev, cudaStream_t s, bool have_occupancy);
