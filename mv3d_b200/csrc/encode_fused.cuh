// encode_fused.cuh -- the TILE strategy as ONE persistent kernel (included by encode_kernels.cu).
//
// Why: as four kernels in a row the step is classify (issue-bound, HBM at 25 %) + scan + bin, THEN the tile kernel
// (store-bound, the SMs' issue slots half empty): 17 % of the step leaves HBM idle (profiles/bench_r2d.json). The two
// kinds of work want different resources, so they are interleaved on the same SMs instead: every sweep still goes
// classify -> scan -> bin -> tile build, but the tile build of sweep s-3 runs beside the bin pass of sweep s-2, the
// scan of sweep s-1 and the classify pass of sweep s.
//
// How: the work of a call is cut into ITEMS -- a chunk of points to classify (C), one sweep's histogram to scan (S), a
// chunk of points to bin (B), a chunk of front-view pixels to finish (F), one output tile to build and store (T) --
// laid out in one global order: stage s = { C(s), S(s-1), B(s-2) + F(s-2), T(s-3) }, point items spread evenly between
// the tile items of the stage. Persistent CTAs (as many as fit the GPU, like the tile kernel's) take the next item
// with an atomic ticket. An item that needs another phase's output waits for that sweep's completion counter
// (acquire load; the producers bump it with a release). Tickets are handed out in order and a CTA only ever waits for
// items with LOWER tickets, which are held by CTAs that are running, so the scheme cannot deadlock whatever the
// residency. The per-item arithmetic is the same code the four kernels run (classify_as / quantise / the shared-memory
// reduction of tile_body), so the result is bit-identical; MV3D_FUSED=0 falls back to the four launches.
//
// Data produced by another CTA of the SAME kernel is read with ld.global.cg (L2), never through the non-coherent path.
#pragma once

namespace mv3d {

#ifndef MV3D_FUSED_CLASSIFY_CHUNK
#define MV3D_FUSED_CLASSIFY_CHUNK 1024
#endif
#ifndef MV3D_FUSED_BIN_CHUNK
#define MV3D_FUSED_BIN_CHUNK 2048
#endif
constexpr int kFusedClassifyChunk = MV3D_FUSED_CLASSIFY_CHUNK;  // points per C item (rounds of 256 threads x kPointItems)
constexpr int kFusedBinChunk = MV3D_FUSED_BIN_CHUNK;            // points per B item
constexpr int kFusedFrontChunk = 8192;     // pixels per F item
constexpr int kFusedLag = 3;               // stages between a sweep's classify pass and its tiles

// sync words (zeroed before every launch): [0] ticket, then per sweep: classify items done, scan done, bin items done
__host__ __device__ inline int fused_sync_words(int batch) { return 1 + 3 * batch; }

__device__ __forceinline__ unsigned ld_acquire(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// Block until *counter >= need (one thread polls, the CTA follows through the barrier).
__device__ __forceinline__ void fused_wait(const unsigned *counter, unsigned need)
{
    if (threadIdx.x == 0) {
        while (ld_acquire(counter) < need) __nanosleep(64);
    }
    __syncthreads();
}
// Publish this CTA's writes, then count the item as done.
__device__ __forceinline__ void fused_signal(unsigned *counter)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        atomicAdd(counter, 1u);
    }
}

// ---- C item: classify points [i_begin, i_end) of sweep f (the body of classify_kernel) -------------------------------
template <bool kPy>
__device__ __forceinline__ void fused_classify(const float4 *__restrict__ pts, const FrameOffsets &offs, const ViewParams &p,
                                               uint2 *__restrict__ keys, unsigned *__restrict__ tile_count,
                                               unsigned *__restrict__ fv_winner, int f, int i_begin, int i_end)
{
    const int base = offs.off[f];
    const int n = min(offs.off[f + 1] - base, i_end);
    const unsigned lane = threadIdx.x & 31u;
    unsigned *counts = tile_count + (size_t)f * (p.tiles_per_frame + 1);
    unsigned *winners = fv_winner ? fv_winner + (size_t)f * p.front_pixels : nullptr;
    const int stride = blockDim.x;
    const int warp_first = i_begin + (int)threadIdx.x - (int)lane;
    for (int i0 = warp_first; i0 < n; i0 += stride * kPointItems) {
        float4 pt[kPointItems];
        int idx[kPointItems];
#pragma unroll
        for (int k = 0; k < kPointItems; k++) {
            idx[k] = i0 + k * stride + (int)lane;
            if (idx[k] < n) pt[k] = load_point(pts + base + idx[k]);
        }
        uint32_t cell[kPointItems], tile[kPointItems];
#pragma unroll
        for (int k = 0; k < kPointItems; k++) {
            cell[k] = kInvalidKey;
            tile[k] = kInvalidKey;
            if (idx[k] < n) {
                const CellHit c = classify_as<kPy>(pt[k], p);
                if (c.keep) {
                    cell[k] = c.cell;
                    if (!kPy && winners) {
                        const int pix = front_pixel(pt[k], p);
                        if (pix >= 0) atomicMax(winners + pix, (unsigned)idx[k] + 1u);
                    }
                }
            }
            if (cell[k] != kInvalidKey) tile[k] = tile_of(cell[k], p);
        }
        unsigned peers[kPointItems], rank[kPointItems];
        int leader[kPointItems];
#pragma unroll
        for (int k = 0; k < kPointItems; k++) {
            if (i0 + k * stride >= n) { peers[k] = 0u; leader[k] = 0; rank[k] = 0u; continue; }  // warp-uniform
            peers[k] = __match_any_sync(0xFFFFFFFFu, tile[k]);
            leader[k] = __ffs(peers[k]) - 1;
            rank[k] = 0;
            if (tile[k] != kInvalidKey && (int)lane == leader[k])
                rank[k] = atomicAdd(counts + tile[k], (unsigned)__popc(peers[k]));
        }
#pragma unroll
        for (int k = 0; k < kPointItems; k++) {
            if (i0 + k * stride >= n) continue;
            unsigned r = __shfl_sync(0xFFFFFFFFu, rank[k], leader[k]);
            r += (unsigned)__popc(peers[k] & ((1u << lane) - 1u));
            if (idx[k] < n) keys[base + idx[k]] = make_uint2(cell[k], r);
        }
    }
}

// ---- S item: exclusive scan of sweep f's tile histogram, clearing it (the body of tile_scan_kernel, any CTA size) ----
__device__ __forceinline__ void fused_scan(unsigned *__restrict__ tile_count, unsigned *__restrict__ tile_off,
                                           int tiles_per_frame, int f, unsigned *warp_sums /* shared [33] */)
{
    unsigned *cnt = tile_count + (size_t)f * (tiles_per_frame + 1);
    unsigned *off = tile_off + (size_t)f * (tiles_per_frame + 1);
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    constexpr int kItems = 4;
    if (threadIdx.x == 0) warp_sums[32] = 0u;  // running total
    __syncthreads();
    for (int t0 = 0; t0 < tiles_per_frame; t0 += (int)blockDim.x * kItems) {
        const int first = t0 + (int)threadIdx.x * kItems;
        unsigned v[kItems], sum = 0;
#pragma unroll
        for (int j = 0; j < kItems; j++) {
            v[j] = 0;
            if (first + j < tiles_per_frame) { v[j] = __ldcg(cnt + first + j); cnt[first + j] = 0; }
            sum += v[j];
        }
        unsigned incl = sum;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned up = __shfl_up_sync(0xFFFFFFFFu, incl, d);
            if ((int)lane >= d) incl += up;
        }
        if (lane == 31) warp_sums[warp] = incl;
        __syncthreads();
        if (warp == 0) {
            const unsigned w = (lane < nwarps) ? warp_sums[lane] : 0u;
            unsigned wincl = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned up = __shfl_up_sync(0xFFFFFFFFu, wincl, d);
                if ((int)lane >= d) wincl += up;
            }
            if (lane < nwarps) warp_sums[lane] = wincl - w;
        }
        __syncthreads();
        unsigned excl = warp_sums[32] + warp_sums[warp] + incl - sum;
#pragma unroll
        for (int j = 0; j < kItems; j++) {
            if (first + j < tiles_per_frame) off[first + j] = excl;
            excl += v[j];
        }
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) warp_sums[32] = excl;
        __syncthreads();
    }
    if (threadIdx.x == 0) off[tiles_per_frame] = warp_sums[32];
}

// ---- B item: move the kept points [i_begin, i_end) of sweep f into their tiles' bins (the body of bin_kernel) --------
template <bool kPy>
__device__ __forceinline__ void fused_bin(const float4 *__restrict__ pts, const FrameOffsets &offs, const ViewParams &p,
                                          const uint2 *__restrict__ keys, const unsigned *__restrict__ tile_off,
                                          uint4 *__restrict__ binned, int f, int i_begin, int i_end)
{
    const int base = offs.off[f];
    const int n = min(offs.off[f + 1] - base, i_end);
    const unsigned *off = tile_off + (size_t)f * (p.tiles_per_frame + 1);
    const int stride = blockDim.x;
    for (int i0 = i_begin + (int)threadIdx.x; i0 < n; i0 += stride * kPointItems) {
        uint2 k[kPointItems];
        float4 pt[kPointItems];
        unsigned start[kPointItems];
#pragma unroll
        for (int j = 0; j < kPointItems; j++) {
            const int i = i0 + j * stride;
            k[j] = make_uint2(kInvalidKey, 0u);
            if (i < n) k[j] = __ldcg(keys + base + i);
        }
#pragma unroll
        for (int j = 0; j < kPointItems; j++) {
            start[j] = 0u;
            if (k[j].x != kInvalidKey) {
                pt[j] = load_point(pts + base + i0 + j * stride);
                start[j] = __ldcg(off + tile_of(k[j].x, p));
            }
        }
#pragma unroll
        for (int j = 0; j < kPointItems; j++) {
            if (k[j].x == kInvalidKey) continue;
            const int i = i0 + j * stride;
            const uint32_t tile = tile_of(k[j].x, p);
            const uint32_t local = k[j].x - tile * (uint32_t)p.tile_cells;
            uint32_t zfield;
            float key;
            if (kPy) {
                const CellHit c = py_classify(pt[j], p);
                key = c.h;
                zfield = (c.slice < p.zn ? (uint32_t)c.slice : kNoSlice) | ((c.lower && c.slice <= p.zn) ? kLowerFlag : 0u);
            } else {
                zfield = (uint32_t)quantise(pt[j].z, p.z_min, p.z_div, p.z_inv, p.quant_floor);
                key = __fsub_rn(pt[j].z, p.z_min);
            }
            binned[(size_t)base + start[j] + k[j].y] =
                make_uint4(local | (zfield << 16), __float_as_uint(key), (uint32_t)i, __float_as_uint(pt[j].w));
        }
    }
}

// ---- F item: front-view pixels [pix_begin, pix_end) of sweep f from the elected winners; clears the winners ---------
__device__ __forceinline__ void fused_front(const float4 *__restrict__ pts, const FrameOffsets &offs, const ViewParams &p,
                                            unsigned *__restrict__ fv_winner, float *__restrict__ front, int f,
                                            int pix_begin, int pix_end)
{
    unsigned *winners = fv_winner + (size_t)f * p.front_pixels;
    float *out = front + (size_t)f * p.front_frame_floats;
    const int end = min(pix_end, p.front_pixels);
    for (int pix = pix_begin + (int)threadIdx.x; pix < end; pix += blockDim.x) {
        const unsigned w = __ldcg(winners + pix);
        float v0 = 0.0f, v1 = 0.0f, v2 = 0.0f;
        if (w) {
            winners[pix] = 0u;
            const float4 pt = pts[offs.off[f] + (int)(w - 1u)];
            front_values(pt, p, v0, v1, v2);
        }
        float *o = out + (size_t)pix * p.fn;
        o[0] = v0; o[1] = v1; o[2] = v2;
        for (int c = 3; c < p.fn; c++) o[c] = 0.0f;
    }
}

// ---- T item: build and store output tile tl of sweep f (one iteration of tile_body's loop, dense bulk-store form) ----
// img / key / dirty: the CTA's shared-memory tile image (all-zero between tiles), winner keys and touched-cell list.
// The head of a tile's bin, fetched one item ahead so that its L2 round trips are off the tile's critical path.
struct TileHead {
    unsigned beg;
    int n;
    uint4 q0;     // record `threadIdx.x` of the bin (if there is one)
    bool valid;
};
__device__ __forceinline__ TileHead fused_tile_head(const uint4 *__restrict__ binned, const unsigned *__restrict__ tile_off,
                                                    const FrameOffsets &offs, const ViewParams &p, int f, int tl)
{
    TileHead h;
    const unsigned *off = tile_off + (size_t)f * (p.tiles_per_frame + 1) + tl;
    h.beg = __ldcg(off);
    h.n = (int)(__ldcg(off + 1) - h.beg);
    h.q0 = make_uint4(0, 0, 0, 0);
    if ((int)threadIdx.x < h.n) h.q0 = __ldcg(binned + (size_t)offs.off[f] + h.beg + threadIdx.x);
    h.valid = true;
    return h;
}

template <bool kPy>
__device__ __forceinline__ void fused_tile(const uint4 *__restrict__ binned, const TileHead &head,
                                           const FrameOffsets &offs, const ViewParams &p, const float *__restrict__ lut,
                                           float *__restrict__ top, int f, int tl, unsigned *img,
                                           unsigned long long *key, unsigned *dirty, unsigned *ndirty_p)
{
    constexpr int kInFlight = kPy ? 1 : kRecordsInFlight;
    constexpr int kWarps = kTileThreads / 32;
    const int T = p.tile_cells, planes = p.planes, zn = p.zn;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n0 = head.n;
    const uint4 *rec0 = binned + (size_t)offs.off[f] + head.beg;

    const long long cells_total = (long long)p.xn * p.yn;
    const long long cell0 = (long long)tl * T;
    const int ncell = (int)min((long long)T, cells_total - cell0);
    const uint32_t bytes = (uint32_t)ncell * (uint32_t)planes * 4u;
    float *dst = top + (size_t)f * p.top_frame_floats + (size_t)cell0 * planes;
    const bool bulk = ((reinterpret_cast<uintptr_t>(dst) | (uintptr_t)bytes) & 15u) == 0;

    if (n0 == 0) {  // empty tile: stream the all-zero image; nothing is modified, so no barrier
        if (bulk) {
            if (tid == 0) { bulk_store_hint(dst, img, bytes, l2_policy_evict_first()); bulk_commit(); }
        } else {
            for (int w = tid; w < ncell * planes; w += kTileThreads) dst[w] = 0.0f;
        }
        return;
    }
    if (tid == 0) bulk_wait_read_all();  // earlier (empty-tile) bulk stores may still be reading the image
    __syncthreads();

    auto accumulate = [&](const uint4 &q) {
        const unsigned local = q.x & 0xFFFFu, Z = q.x >> 16;
        unsigned *cellw = img + local * planes;
        if (atomicAdd(cellw + zn + 1, 1u) == 0u) dirty[atomicAdd(ndirty_p, 1u)] = local;
        if ((int)q.y > 0) {  // h > 0: the reference's maps start at 0 and update on strict '>'
            if (kPy) {
                const unsigned zs = Z & (kLowerFlag - 1u);
                if (zs != kNoSlice) atomicMax(cellw + zs, q.y);
                if (Z & kLowerFlag) {
                    const unsigned below = (zs == kNoSlice) ? (unsigned)zn - 1u : zs - 1u;
                    if (below < (unsigned)zn) atomicMax(cellw + below, q.y);
                }
            } else {
                atomicMax(cellw + Z, q.y);
            }
            atomicMax(key + local, ((unsigned long long)q.y << 32) | (unsigned long long)(0xFFFFFFFFu - q.z));
        }
    };
    const uint4 q0 = head.q0;
    if (tid < n0) accumulate(q0);
    for (int r = tid + kTileThreads; r < n0; r += kInFlight * kTileThreads) {
        uint4 q[kInFlight];
#pragma unroll
        for (int j = 0; j < kInFlight; j++)
            if (r + j * kTileThreads < n0) q[j] = __ldcg(rec0 + r + j * kTileThreads);
#pragma unroll
        for (int j = 0; j < kInFlight; j++)
            if (r + j * kTileThreads < n0) accumulate(q[j]);
    }
    __syncthreads();

    const int nd = (int)*ndirty_p;
    auto claim = [&](const uint4 &q) {  // intensity of each cell's winner
        const unsigned local = q.x & 0xFFFFu;
        if ((int)q.y > 0 && key[local] == (((unsigned long long)q.y << 32) | (unsigned long long)(0xFFFFFFFFu - q.z)))
            img[local * planes + zn] = q.w;
    };
    if (tid < n0) claim(q0);
    for (int r = tid + kTileThreads; r < n0; r += kInFlight * kTileThreads) {
        uint4 q[kInFlight];
#pragma unroll
        for (int j = 0; j < kInFlight; j++)
            if (r + j * kTileThreads < n0) q[j] = __ldcg(rec0 + r + j * kTileThreads);
#pragma unroll
        for (int j = 0; j < kInFlight; j++)
            if (r + j * kTileThreads < n0) claim(q[j]);
    }
    for (int d = warp; d < nd; d += kWarps) {  // one warp per touched cell, lanes over planes
        unsigned *cellw = img + dirty[d] * planes;
        for (int k = lane; k < planes; k += 32) {
            if (k < zn) {
                const unsigned raw = cellw[k];
                if (raw) cellw[k] = __float_as_uint(kPy ? __fsub_rn(__uint_as_float(raw), (float)k)
                                                         : slice_value(__uint_as_float(raw), p.z_div, k));
            } else if (k == zn + 1) {
                const unsigned c = cellw[k];
                cellw[k] = __float_as_uint(__ldg(lut + (c < 31u ? c : 31u)));
            }
        }
    }
    fence_proxy_async_smem();
    __syncthreads();
    if (bulk) {
        if (tid == 0) { bulk_store_hint(dst, img, bytes, l2_policy_evict_first()); bulk_commit(); bulk_wait_read_all(); }
    } else {
        for (int w = tid; w < ncell * planes; w += kTileThreads) dst[w] = __uint_as_float(img[w]);
    }
    __syncthreads();
    for (int d = warp; d < nd; d += kWarps) {
        const unsigned local = dirty[d];
        for (int k = lane; k < planes; k += 32) img[local * planes + k] = 0u;
        if (lane == 0) key[local] = 0ull;
    }
    if (tid == 0) *ndirty_p = 0;
    fence_proxy_async_smem();
    __syncthreads();
}

// What one launch covers, by value in the kernel parameter space.
struct FusedPlan {
    int batch;
    int n_classify, n_bin, n_front;  // C / B / F items per sweep (from the longest sweep of the call)
    int n_point;                     // point items per stage: n_classify + 1 (scan) + n_bin + n_front
    int per_stage;                   // n_point + tiles_per_frame
    int stages;                      // batch + kFusedLag
};

template <bool kPy>
__global__ void __launch_bounds__(kTileThreads, MV3D_TILE_MIN_BLOCKS)
encode_fused_kernel(const float4 *__restrict__ pts, const __grid_constant__ FrameOffsets offs,
                    const __grid_constant__ ViewParams p, const __grid_constant__ FusedPlan plan,
                    uint2 *__restrict__ keys, unsigned *__restrict__ tile_count, unsigned *__restrict__ tile_off,
                    uint4 *__restrict__ binned, unsigned *__restrict__ fv_winner, float *__restrict__ front,
                    const float *__restrict__ lut, float *__restrict__ top, unsigned *__restrict__ sync)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int T = p.tile_cells, planes = p.planes;
    unsigned *img = reinterpret_cast<unsigned *>(smem_raw);
    const size_t img_bytes = ((size_t)T * planes * 4 + 15) & ~(size_t)15;
    unsigned long long *key = reinterpret_cast<unsigned long long *>(smem_raw + img_bytes);
    unsigned *dirty = reinterpret_cast<unsigned *>(key + T);
    // tk[parity] = {this iteration's ticket, the next one}: thread 0 writes the other parity's pair during an
    // iteration, so ONE barrier per item separates writers from readers. ready_tiles: sweeps below it are fully binned
    // (checked once per sweep and CTA, not once per tile).
    __shared__ unsigned ndirty, tk[2][2], warp_sums[33];
    __shared__ int ready_tiles;
    const int tid = threadIdx.x;

    for (int w = tid; w < T * planes; w += kTileThreads) img[w] = 0u;
    for (int w = tid; w < T; w += kTileThreads) key[w] = 0ull;
    if (tid == 0) {
        ndirty = 0;
        ready_tiles = 0;
        tk[0][0] = atomicAdd(sync, 1u);
        tk[0][1] = atomicAdd(sync, 1u);
    }
    fence_proxy_async_smem();
    __syncthreads();

    unsigned *c_done = sync + 1, *s_done = sync + 1 + plan.batch, *b_done = sync + 1 + 2 * plan.batch;
    const unsigned total = (unsigned)plan.stages * (unsigned)plan.per_stage;
    const int tpf = p.tiles_per_frame;
    // ticket -> (is a tile item, sweep, index); point items are spread evenly between the tile items of a stage
    auto decode = [&](unsigned ticket, bool &is_tile, int &stage, int &index) {
        stage = (int)(ticket / (unsigned)plan.per_stage);
        const int slot = (int)(ticket - (unsigned)stage * (unsigned)plan.per_stage);
        const int before = (int)(((long long)slot * plan.n_point) / plan.per_stage);
        is_tile = (int)(((long long)(slot + 1) * plan.n_point) / plan.per_stage) == before;
        index = is_tile ? slot - before : before;
    };
    TileHead have;
    have.valid = false;
    for (int it = 0;; it++) {
        const int par = it & 1;
        const unsigned ticket = tk[par][0], ticket_next = tk[par][1];
        if (ticket >= total) break;
        unsigned ticket_after = 0u;  // requested now, needed at the end of the item: stays in a register meanwhile
        if (tid == 0) ticket_after = atomicAdd(sync, 1u);
        const int ready = ready_tiles;
        bool is_tile, next_is_tile;
        int stage, index, next_stage, next_index;
        decode(ticket, is_tile, stage, index);
        TileHead ahead;
        ahead.valid = false;
        if (ticket_next < total) {
            decode(ticket_next, next_is_tile, next_stage, next_index);
            const int nf = next_stage - kFusedLag;
            if (next_is_tile && nf >= 0 && nf < ready && next_index < tpf)
                ahead = fused_tile_head(binned, tile_off, offs, p, nf, next_index);
        }
        if (is_tile) {
            const int f = stage - kFusedLag;
            if (f >= 0 && f < plan.batch && index < tpf) {
                if (f >= ready) {  // first tile of this sweep on this CTA: wait until all its points are binned
                    fused_wait(b_done + f, (unsigned)plan.n_bin);
                    if (tid == 0) ready_tiles = f + 1;
                    have.valid = false;
                }
                if (!have.valid) have = fused_tile_head(binned, tile_off, offs, p, f, index);
                fused_tile<kPy>(binned, have, offs, p, lut, top, f, index, img, key, dirty, &ndirty);
            }
        } else if (index < plan.n_classify) {
            if (stage < plan.batch) {
                fused_classify<kPy>(pts, offs, p, keys, tile_count, fv_winner, stage, index * kFusedClassifyChunk,
                                    (index + 1) * kFusedClassifyChunk);
                fused_signal(c_done + stage);
            }
        } else if (index == plan.n_classify) {
            const int f = stage - 1;
            if (f >= 0 && f < plan.batch) {
                fused_wait(c_done + f, (unsigned)plan.n_classify);
                fused_scan(tile_count, tile_off, tpf, f, warp_sums);
                fused_signal(s_done + f);
            }
        } else if (index <= plan.n_classify + plan.n_bin) {
            const int f = stage - 2, c = index - plan.n_classify - 1;
            if (f >= 0 && f < plan.batch) {
                fused_wait(s_done + f, 1u);
                fused_bin<kPy>(pts, offs, p, keys, tile_off, binned, f, c * kFusedBinChunk, (c + 1) * kFusedBinChunk);
                fused_signal(b_done + f);
            }
        } else {
            const int f = stage - 2, c = index - plan.n_classify - 1 - plan.n_bin;
            if (f >= 0 && f < plan.batch) {
                fused_wait(c_done + f, (unsigned)plan.n_classify);
                fused_front(pts, offs, p, fv_winner, front, f, c * kFusedFrontChunk, (c + 1) * kFusedFrontChunk);
            }
        }
        have = ahead;
        if (tid == 0) { tk[par ^ 1][0] = ticket_next; tk[par ^ 1][1] = ticket_after; }
        __syncthreads();  // the item is over for the whole CTA, and the next pair of tickets is published
    }
    if (tid == 0) bulk_wait_read_all();  // shared memory must outlive the last bulk read
}

} // namespace mv3d
