// encode_kernels.cu -- sm_100a kernels of the LiDAR view encoder and their launchers.
//
// Two strategies for the per-cell reduction of the bird's-eye view (BEV), chosen per parameter set:
//
//  TILE (default)  classify -> scan -> bin -> tile_build
//      The output [X][Y][Z+2] is cell-major, so any run of consecutive cells is one contiguous chunk of
//      memory. Points are binned by output tile (a counting sort keyed on tile id; the rank inside the bin
//      comes from the same atomic that builds the histogram). One CTA then owns one tile at a time: it
//      reduces the tile's points in shared memory with order-preserving integer atomics, finalises the
//      touched cells in place, and writes the tile -- zeros included -- with ONE bulk async (TMA) store.
//      Every output byte is written exactly once and never read; HBM traffic = the algorithmic bytes.
//
//  PATCH           copy-engine zero-fill (side stream)  ||  classify -> scan -> bin ; then tile_patch per chunk of sweeps
//      The zero-fill of the output needs no point data and no SM at all when the copy engine does it
//      (cudaMemsetAsync: 7.0-7.3 TB/s), so it runs beside the point passes; the tile reduction then writes only
//      the touched cells into the (already zero) output. Kept as the measured comparison point of DESIGN.md 5.1.
//
//  ATOMIC          fill_zero -> atomic_scatter -> atomic_finalize
//      Zero-fill the output, reduce with global atomics straight into the output cells (tagged raw values),
//      then a second pass over the points converts the touched slots. No binning; relies on L2 atomics.
//
// The front view (FV) is a per-pixel "last point in input order wins" image: classify/scatter records the
// winning point index per pixel with atomicMax, front_finalize recomputes the three channel values from the
// winner and writes the whole image (zeros included) once.
#include "encode_kernels.h"
#include "view_math.cuh"
#include "selfcheck.cuh"

#include <mutex>

namespace mv3d {

constexpr int kPointThreads = 256;
constexpr int kTileThreads = 256;
#ifndef MV3D_TILE_MIN_BLOCKS
#define MV3D_TILE_MIN_BLOCKS 5  // register cap of the tile kernels: five CTAs of 256 threads per SM (48 registers)
#endif

// ------------------------------------------------------------------------------------------------
// small PTX wrappers
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_addr(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}
// Make generic-proxy writes to shared memory visible to the async proxy (TMA) before a bulk store reads them.
__device__ __forceinline__ void fence_proxy_async_smem()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
// Bulk async copy shared -> global (TMA, non-tensor form). bytes % 16 == 0, both addresses 16-byte aligned.
__device__ __forceinline__ void bulk_store(void *gdst, const void *ssrc, uint32_t bytes)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 :: "l"(gdst), "r"(smem_addr(ssrc)), "r"(bytes) : "memory");
}
// Same, with an L2 eviction policy: the zero-fill stream is written once and never read by this library, so it
// should leave L2 first and not push the point/bin scratch out.
__device__ __forceinline__ uint64_t l2_policy_evict_first()
{
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    return pol;
}
__device__ __forceinline__ void bulk_store_hint(void *gdst, const void *ssrc, uint32_t bytes, uint64_t policy)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;"
                 :: "l"(gdst), "r"(smem_addr(ssrc)), "r"(bytes), "l"(policy) : "memory");
}
// Wait until all committed bulk stores of this thread are complete (writes performed, not only the smem read).
__device__ __forceinline__ void bulk_wait_all()
{
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}
// Wait until at most N committed bulk stores of this thread are still pending (full completion).
template <int N>
__device__ __forceinline__ void bulk_wait_pending()
{
    asm volatile("cp.async.bulk.wait_group %0;" :: "n"(N) : "memory");
}
__device__ __forceinline__ void bulk_commit()
{
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
// Wait until all committed bulk stores of this thread have finished READING shared memory.
__device__ __forceinline__ void bulk_wait_read_all()
{
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
}
// How the tile kernel writes a finished tile: one bulk store of the whole image, with an L2 evict-first policy -- the
// output is written once and never read back here, so it should not push the bin records (read by this same kernel)
// out of L2 -- or, when the copy engine has zero-filled the output already (PATCH), the touched cells only.
// (Round 1 also measured plain / split / paced bulk stores and LSU copies of the image: tools/experiments/.)
enum TileStore { kStoreBulk = 0, kStorePatch = 1 };
// Programmatic dependent launch (griddepcontrol): a kernel launched with the programmatic-serialization attribute may
// start while its predecessor in the stream is still draining; pdl_wait() blocks until that predecessor has completed
// and its writes are visible, pdl_trigger() lets the successor start launching once every CTA of this grid has got
// here (or exited). Both are no-ops when the launch carries no such dependency.
__device__ __forceinline__ void pdl_wait()
{
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
__device__ __forceinline__ void pdl_trigger()
{
    asm volatile("griddepcontrol.launch_dependents;");
}
__device__ __forceinline__ void store_zero16(float4 *p)
{
    asm volatile("st.global.cs.v4.f32 [%0], {%1,%1,%1,%1};" :: "l"(p), "f"(0.0f) : "memory");
}

// ------------------------------------------------------------------------------------------------
// TILE strategy, pass 1: classify every point, histogram tiles, elect front-view winners.
//   keys[g] = {output cell, rank inside its tile's bin}  ({kInvalidKey, 0} when the point is rejected)
// grid = (blocks, batch); each CTA strides over one sweep's points with coalesced 16-byte loads.
// ------------------------------------------------------------------------------------------------
#ifndef MV3D_POINT_ITEMS
#define MV3D_POINT_ITEMS 2
#endif
constexpr int kPointItems = MV3D_POINT_ITEMS;  // points per thread and round: their loads and atomics are in flight together
#ifndef MV3D_CLASSIFY_MIN_BLOCKS
#define MV3D_CLASSIFY_MIN_BLOCKS 8  // 32 registers (24 bytes spilled): 64 warps per SM instead of 40; measured 0.096 -> 0.078 ms per 64 sweeps
#endif
template <bool kPy>
__global__ void __launch_bounds__(kPointThreads, MV3D_CLASSIFY_MIN_BLOCKS)
classify_kernel(const float4 *__restrict__ pts, const __grid_constant__ FrameOffsets offs,
                const __grid_constant__ ViewParams p, uint2 *__restrict__ keys, unsigned *__restrict__ tile_count,
                unsigned *__restrict__ fv_winner)
{
    pdl_trigger();
    const int f = blockIdx.y;
    const int base = offs.off[f];
    const int n = offs.off[f + 1] - base;
    const unsigned lane = threadIdx.x & 31u;
    unsigned *counts = tile_count + (size_t)f * (p.tiles_per_frame + 1);
    unsigned *winners = fv_winner + (size_t)f * p.front_pixels;

    // warp-uniform trip count so the match/shuffle below is convergent; a round covers kPointItems consecutive
    // chunks of gridDim.x * blockDim.x points, so every load stays coalesced
    const int stride = gridDim.x * blockDim.x;
    const int first = blockIdx.x * blockDim.x + threadIdx.x;
    const int warp_first = first - (int)lane;
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
                    if (!kPy && p.want_front) {
                        const int pix = front_pixel(pt[k], p);
                        MV3D_CHECK(pix < p.front_pixels, MV3D_BAD_PIXEL);
                        if (pix >= 0) atomicMax(winners + pix, (unsigned)idx[k] + 1u);
                    }
                }
            }
            if (cell[k] != kInvalidKey) tile[k] = tile_of(cell[k], p);
            MV3D_CHECK(cell[k] == kInvalidKey || tile[k] < (uint32_t)p.tiles_per_frame, MV3D_BAD_TILE);
        }
        if (p.want_top) {
            // Warp-aggregated histogram: lanes that hit the same tile share one atomicAdd. All the atomics of a round
            // are issued before the first result is consumed.
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
}

// ------------------------------------------------------------------------------------------------
// TILE strategy, pass 2: exclusive scan of one sweep's tile histogram (one CTA per sweep).
// Writes tiles_per_frame+1 offsets and clears the histogram for the next call (self-cleaning workspace).
// ------------------------------------------------------------------------------------------------
constexpr int kScanThreads = 1024;
// kScanItems consecutive entries per thread and round (2 for the tile histograms, 32 for the 150k-pixel histogram of
// the Python front view: one CTA per sweep, so the number of rounds is what costs).
template <int kScanItems>
__global__ void __launch_bounds__(kScanThreads)
tile_scan_kernel(unsigned *__restrict__ tile_count, unsigned *__restrict__ tile_off, int tiles_per_frame)
{
    __shared__ unsigned warp_sums[kScanThreads / 32];
    __shared__ unsigned carry;
    unsigned *cnt = tile_count + (size_t)blockIdx.x * (tiles_per_frame + 1);
    unsigned *off = tile_off + (size_t)blockIdx.x * (tiles_per_frame + 1);
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    pdl_trigger();
    pdl_wait();  // the histogram is classify's output
    __syncthreads();
    for (int t0 = 0; t0 < tiles_per_frame; t0 += kScanThreads * kScanItems) {
        const int first = t0 + (int)threadIdx.x * kScanItems;
        unsigned v[kScanItems];
        unsigned sum = 0;
#pragma unroll
        for (int j = 0; j < kScanItems; j++) {
            v[j] = 0;
            if (first + j < tiles_per_frame) { v[j] = cnt[first + j]; cnt[first + j] = 0; }
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
            unsigned w = (lane < kScanThreads / 32) ? warp_sums[lane] : 0u;
            unsigned wincl = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const unsigned up = __shfl_up_sync(0xFFFFFFFFu, wincl, d);
                if ((int)lane >= d) wincl += up;
            }
            if (lane < kScanThreads / 32) warp_sums[lane] = wincl - w;  // exclusive prefix of the warp totals
        }
        __syncthreads();
        unsigned excl = carry + warp_sums[warp] + incl - sum;  // entries before this thread's first one
#pragma unroll
        for (int j = 0; j < kScanItems; j++) {
            if (first + j < tiles_per_frame) off[first + j] = excl;
            excl += v[j];
        }
        __syncthreads();
        if (threadIdx.x == kScanThreads - 1) carry = excl;
        __syncthreads();
    }
    if (threadIdx.x == 0) off[tiles_per_frame] = carry;
}

// ------------------------------------------------------------------------------------------------
// Front view, second half: write the whole [R][C][F] image of sweep f from the per-pixel winners classify elected,
// and clear the winners (self-cleaning workspace). Runs as extra CTAs of the bin pass (both only need classify's
// output), or as a kernel of its own where there is no bin pass.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void front_finalize_body(const float4 *__restrict__ pts, const FrameOffsets &offs,
                                                    const ViewParams &p, unsigned *__restrict__ fv_winner,
                                                    float *__restrict__ front, int f, int block, int blocks)
{
    unsigned *winners = fv_winner + (size_t)f * p.front_pixels;
    float *out = front + (size_t)f * p.front_frame_floats;
    for (int pix = block * blockDim.x + threadIdx.x; pix < p.front_pixels; pix += blocks * blockDim.x) {
        const unsigned w = winners[pix];
        float v0 = 0.0f, v1 = 0.0f, v2 = 0.0f;
        if (w) {
            winners[pix] = 0u;
            MV3D_CHECK((int)(w - 1u) < offs.off[f + 1] - offs.off[f], MV3D_BAD_WINNER);
            const float4 pt = pts[offs.off[f] + (int)(w - 1u)];
            front_values(pt, p, v0, v1, v2);
        }
        float *o = out + (size_t)pix * p.fn;
        o[0] = v0; o[1] = v1; o[2] = v2;
        for (int c = 3; c < p.fn; c++) o[c] = 0.0f;
    }
}

// ------------------------------------------------------------------------------------------------
// TILE strategy, pass 3: move each kept point's record into its tile's bin.
//   record = { cell-in-tile | slice << 16,  h bits,  point index,  intensity bits }   (16 bytes)
// Order inside a bin is arbitrary; every order-dependent decision uses the point index carried along.
// The last front_blocks CTAs of every sweep's row of the grid finish the front view instead (front_finalize_body).
// ------------------------------------------------------------------------------------------------
template <bool kPy>
__global__ void __launch_bounds__(kPointThreads)
bin_kernel(const float4 *__restrict__ pts, const __grid_constant__ FrameOffsets offs,
           const __grid_constant__ ViewParams p, const uint2 *__restrict__ keys,
           const unsigned *__restrict__ tile_off, uint4 *__restrict__ binned, unsigned *__restrict__ fv_winner,
           float *__restrict__ front, int front_blocks)
{
    pdl_trigger();
    pdl_wait();  // keys / winners are classify's output, the bin offsets are the scan's
    const int f = blockIdx.y;
    const int bin_blocks = (int)gridDim.x - front_blocks;
    if ((int)blockIdx.x >= bin_blocks) {
        front_finalize_body(pts, offs, p, fv_winner, front, f, (int)blockIdx.x - bin_blocks, front_blocks);
        return;
    }
    const int base = offs.off[f];
    const int n = offs.off[f + 1] - base;
    const unsigned *off = tile_off + (size_t)f * (p.tiles_per_frame + 1);
    const int stride = bin_blocks * blockDim.x;
    for (int i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += stride * kPointItems) {
        uint2 k[kPointItems];
        float4 pt[kPointItems];
        unsigned start[kPointItems];
#pragma unroll
        for (int j = 0; j < kPointItems; j++) {  // the keys of every item first: independent loads
            const int i = i0 + j * stride;
            k[j] = make_uint2(kInvalidKey, 0u);
            if (i < n) k[j] = keys[base + i];
        }
#ifndef MV3D_BIN_LOAD_ALL
        // Rejected points come in runs (the part of every beam's turn that lies outside the box), so fetching only the
        // kept points' 16 bytes spares about half of the point traffic of this pass.
#pragma unroll
        for (int j = 0; j < kPointItems; j++) {
            start[j] = 0u;
            if (k[j].x != kInvalidKey) {
                pt[j] = load_point(pts + base + i0 + j * stride);
                start[j] = __ldg(off + tile_of(k[j].x, p));
            }
        }
#else
#pragma unroll
        for (int j = 0; j < kPointItems; j++)
            if (i0 + j * stride < n) pt[j] = load_point(pts + base + i0 + j * stride);
#pragma unroll
        for (int j = 0; j < kPointItems; j++)
            start[j] = k[j].x != kInvalidKey ? __ldg(off + tile_of(k[j].x, p)) : 0u;
#endif
#pragma unroll
        for (int j = 0; j < kPointItems; j++) {
            if (k[j].x == kInvalidKey) continue;
            const int i = i0 + j * stride;
            const uint32_t tile = tile_of(k[j].x, p);
            const uint32_t local = k[j].x - tile * (uint32_t)p.tile_cells;
            uint32_t zfield;
            float key;
            if (kPy) {
                // Python twin: key = qz; slice field = floor(qz) (kNoSlice when it is past the last slice, such a
                // point still counts and can win the intensity), top bit = "also closes the slice below"
                const CellHit c = py_classify(pt[j], p);
                key = c.h;
                zfield = (c.slice < p.zn ? (uint32_t)c.slice : kNoSlice) | ((c.lower && c.slice <= p.zn) ? kLowerFlag : 0u);
            } else {
                zfield = (uint32_t)quantise(pt[j].z, p.z_min, p.z_div, p.z_inv, p.quant_floor);
                key = __fsub_rn(pt[j].z, p.z_min);
            }
            MV3D_CHECK(tile < (uint32_t)p.tiles_per_frame, MV3D_BAD_TILE);
            MV3D_CHECK(start[j] + k[j].y < (unsigned)n, MV3D_BAD_BIN_SLOT);
            MV3D_CHECK(local < (uint32_t)p.tile_cells && local < 65536u, MV3D_BAD_LOCAL_CELL);
            binned[(size_t)base + start[j] + k[j].y] =
                make_uint4(local | (zfield << 16), __float_as_uint(key), (uint32_t)i, __float_as_uint(pt[j].w));
        }
    }
}

// ------------------------------------------------------------------------------------------------
// TILE strategy, pass 4: build and store output tiles. Persistent CTAs, tile t -> CTA t mod grid.
//
// Shared memory per CTA: the tile image (tile_cells * planes words, kept all-zero between tiles), one
// 64-bit (h, ~index) key per cell, and the list of cells touched by the current tile.
//   accumulate : slot[k] = max h bits (h > 0 orders like an unsigned int); count++ in the density slot;
//                key = max (h, ~index) -> highest point, earliest in input order on ties
//   finalise   : touched cells only -- heights to h/z_DIV - k, count to the libm-built density table,
//                intensity of the key's winner
//   store      : one cp.async.bulk shared->global for the whole tile (plain stores if misaligned)
//   clean      : re-zero only the touched cells, after the bulk store has finished reading
// Empty tiles (the majority) skip everything but the store, straight from the all-zero image; a non-empty
// tile first waits for those in-flight reads of the image before accumulating into it.
// ------------------------------------------------------------------------------------------------
// kCompact: instead of storing the dense tile, report the touched cells in a per-sweep list
//   compact[f] = one stream of words: per touched cell { cell index, plane masks, non-zero values } (compact_cap words)
//   table[f]   = { first word, touched cells } per tile, then the sweep's word counter (zeroed by the caller)
// A non-empty tile takes the words of its touched cells from the sweep's counter and writes them in ASCENDING cell
// order; the stream is contiguous and holds finished non-zero values only, and the table tells the reader where each
// tile's cells are. This is what the host-buffer path reads back over PCIe instead of the 95 %-zero dense tensor.
// kOccupancy (dense output only): also report, for the empty-anchor filter that consumes the view next
// (anchor_kernels.cu), one bit per cell "touched" -- touched cells carry a density > 0, untouched cells are all
// zero -- into a pre-zeroed per-frame bit map, and raise the frame's flag when an intensity is negative, -0 or NaN
// (the only values of a view that can be; then the filter must take exact sums). The bits of a tile are gathered
// in shared memory; the first and last word of a tile may be shared with its neighbours and go out as atomics.
#ifndef MV3D_RECORDS_IN_FLIGHT
#define MV3D_RECORDS_IN_FLIGHT 2
#endif
constexpr int kRecordsInFlight = MV3D_RECORDS_IN_FLIGHT;  // records of a heavy tile each thread fetches before reducing them
// kInFlight: records of a heavy tile each thread fetches before reducing them. Two pay for the plain C-routine kernel
// (+1.6 %); the occupancy-reporting and Python-twin variants are faster with one (measured: -6.5 % / -0.5 % with two).
template <bool kCompact, bool kOccupancy, bool kPy, int kStore = kStoreBulk,
          int kInFlight = (kCompact || kOccupancy || kPy) ? 1 : kRecordsInFlight>
__device__ __forceinline__ void
tile_body(const uint4 *__restrict__ binned, const unsigned *__restrict__ tile_off, const FrameOffsets &offs,
          const ViewParams &p, const float *__restrict__ lut, float *__restrict__ top, int total_tiles,
          unsigned *__restrict__ compact, unsigned compact_cap, OccupancySink sink = OccupancySink(),
          unsigned *__restrict__ compact_table = nullptr)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int T = p.tile_cells;
    const int planes = p.planes;
    const int zn = p.zn;
    unsigned *img = reinterpret_cast<unsigned *>(smem_raw);
    const size_t img_bytes = ((size_t)T * planes * 4 + 15) & ~(size_t)15;
    unsigned long long *key = reinterpret_cast<unsigned long long *>(smem_raw + img_bytes);
    unsigned *dirty = reinterpret_cast<unsigned *>(key + T);
    unsigned *occ_tile = dirty + T;  // [T/32 + 2], all-zero between tiles (kOccupancy)
    __shared__ unsigned ndirty;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int kWarps = kTileThreads / 32;

    for (int w = tid; w < T * planes; w += kTileThreads) img[w] = 0u;
    for (int w = tid; w < T; w += kTileThreads) key[w] = 0ull;
    if (kOccupancy)
        for (int w = tid; w < T / 32 + 2; w += kTileThreads) occ_tile[w] = 0u;
    if (tid == 0) ndirty = 0;
    fence_proxy_async_smem();
    pdl_trigger();
    pdl_wait();  // everything above overlaps the tail of the bin pass; the bins themselves are its output
    __syncthreads();

    // Software pipeline over this CTA's tiles: the bin offsets of tile t+2 and the first record of tile t+1
    // are requested while tile t is being built, so their L2 latency is off the per-tile critical path.
    const int step = gridDim.x;
    const int tpf = p.tiles_per_frame;
    auto bin_range = [&](int t, const uint4 *&rec, int &n) {
        n = 0;
        rec = binned;
        if (t < total_tiles) {
            const int f = t / tpf, tl = t - f * tpf;
            const unsigned *off = tile_off + (size_t)f * (tpf + 1) + tl;
            const unsigned beg = __ldg(off);
            n = (int)(__ldg(off + 1) - beg);
            MV3D_CHECK(n >= 0 && (long long)beg + n <= (long long)(offs.off[f + 1] - offs.off[f]), MV3D_BAD_BIN_RANGE);
            rec = binned + (size_t)offs.off[f] + beg;
        }
    };
    const uint4 *rec0, *rec1, *rec2;
    int n0, n1, n2;
    bin_range(blockIdx.x, rec0, n0);
    bin_range(blockIdx.x + step, rec1, n1);
    uint4 q0 = make_uint4(0, 0, 0, 0), q1 = q0;
    if (tid < n0) q0 = __ldg(rec0 + tid);

    const long long cells_total = (long long)p.xn * p.yn;
    for (int t = blockIdx.x; t < total_tiles; t += step) {
        bin_range(t + 2 * step, rec2, n2);
        if (tid < n1) q1 = __ldg(rec1 + tid);

        const int f = t / tpf, tl = t - f * tpf;
        const long long cell0 = (long long)tl * T;
        const int ncell = (int)min((long long)T, cells_total - cell0);
        const uint32_t bytes = (uint32_t)ncell * (uint32_t)planes * 4u;
        float *dst = top + (size_t)f * p.top_frame_floats + (size_t)cell0 * planes;
        const bool bulk = ((reinterpret_cast<uintptr_t>(dst) | (uintptr_t)bytes) & 15u) == 0;

        if (kCompact && n0 == 0) {
            // nothing to report for an empty tile but its (empty) range
            if (tid == 0) {
                unsigned *tab = compact_table + (size_t)f * (2u * (unsigned)tpf + 1u) + 2u * (unsigned)tl;
                tab[0] = 0u;
                tab[1] = 0u;
            }
        } else if (n0 == 0) {
            // Empty tile: stream the all-zero image; nothing is modified, so no barrier.
            if (kStore == kStorePatch) {
                // the output is zero already
            } else if (bulk) {
                if (tid == 0) { bulk_store_hint(dst, img, bytes, l2_policy_evict_first()); bulk_commit(); }
            } else {
                for (int w = tid; w < ncell * planes; w += kTileThreads) dst[w] = 0.0f;
            }
        } else {
            // Earlier (empty-tile) bulk stores may still be reading the image: wait before touching it.
            if (!kCompact && kStore == kStoreBulk) {
                if (tid == 0) bulk_wait_read_all();
                __syncthreads();
            }

            // accumulate: the first kTileThreads records are in registers already (prefetched); the rest of a heavy
            // tile (tiles next to the sensor hold hundreds to thousands of records) are fetched four per thread at
            // a time, so that their L2 latency overlaps
            auto accumulate = [&](const uint4 &q) {
                const unsigned local = q.x & 0xFFFFu, Z = q.x >> 16;
                MV3D_CHECK(local < (unsigned)T, MV3D_BAD_LOCAL_CELL);
                MV3D_CHECK(kPy ? ((Z & (kLowerFlag - 1u)) == kNoSlice || (Z & (kLowerFlag - 1u)) < (unsigned)zn) : Z < (unsigned)zn,
                           MV3D_BAD_SLICE);
                unsigned *cellw = img + local * planes;
                if (atomicAdd(cellw + zn + 1, 1u) == 0u) {
                    const unsigned at = atomicAdd(&ndirty, 1u);
                    MV3D_CHECK(at < (unsigned)T, MV3D_BAD_DIRTY);
                    dirty[at] = local;
                }
                if ((int)q.y > 0) {  // h > 0: the reference's maps start at 0 and update on strict '>'
                    if (kPy) {  // slice = floor(qz) if it exists, and the one below when qz is integral
                        const unsigned zs = Z & (kLowerFlag - 1u);
                        if (zs != kNoSlice) atomicMax(cellw + zs, q.y);
                        if (Z & kLowerFlag) {
                            const unsigned below = (zs == kNoSlice) ? (unsigned)zn - 1u : zs - 1u;  // qz == zn closes the last slice
                            if (below < (unsigned)zn) atomicMax(cellw + below, q.y);
                        }
                    } else {
                        atomicMax(cellw + Z, q.y);
                    }
                    atomicMax(key + local, ((unsigned long long)q.y << 32) | (unsigned long long)(0xFFFFFFFFu - q.z));
                }
            };
            if (tid < n0) accumulate(q0);
            for (int r = tid + kTileThreads; r < n0; r += kInFlight * kTileThreads) {
                uint4 q[kInFlight];
#pragma unroll
                for (int j = 0; j < kInFlight; j++)
                    if (r + j * kTileThreads < n0) q[j] = __ldg(rec0 + r + j * kTileThreads);
#pragma unroll
                for (int j = 0; j < kInFlight; j++)
                    if (r + j * kTileThreads < n0) accumulate(q[j]);
            }
            __syncthreads();

            // finalise the touched cells in place
            const int nd = (int)ndirty;
            auto claim = [&](const uint4 &q) {  // intensity of each cell's winner
                const unsigned local = q.x & 0xFFFFu;
                if ((int)q.y > 0 &&
                    key[local] == (((unsigned long long)q.y << 32) | (unsigned long long)(0xFFFFFFFFu - q.z))) {
                    img[local * planes + zn] = q.w;
                    if (kOccupancy && ((q.w >> 31) != 0u || (q.w & 0x7fffffffu) > 0x7f800000u))
                        atomicOr(sink.flags + f, 1u);
                }
            };
            if (tid < n0) claim(q0);
            for (int r = tid + kTileThreads; r < n0; r += kInFlight * kTileThreads) {
                uint4 q[kInFlight];
#pragma unroll
                for (int j = 0; j < kInFlight; j++)
                    if (r + j * kTileThreads < n0) q[j] = __ldg(rec0 + r + j * kTileThreads);
#pragma unroll
                for (int j = 0; j < kInFlight; j++)
                    if (r + j * kTileThreads < n0) claim(q[j]);
            }
            for (int d = warp; d < nd; d += kWarps) {  // one warp per touched cell, lanes over planes
                unsigned *cellw = img + dirty[d] * planes;
                if (kOccupancy && lane == 0) {
                    const unsigned g = (unsigned)cell0 + dirty[d];
                    atomicOr(occ_tile + ((g >> 5) - ((unsigned)cell0 >> 5)), 1u << (g & 31u));
                }
                for (int k = lane; k < planes; k += 32) {
                    if (k < zn) {
                        const unsigned raw = cellw[k];
                        if (raw) cellw[k] = __float_as_uint(kPy ? __fsub_rn(__uint_as_float(raw), (float)k)  // max(qz) - z
                                                                 : slice_value(__uint_as_float(raw), p.z_div, k));
                    } else if (k == zn + 1) {
                        const unsigned c = cellw[k];
                        cellw[k] = __float_as_uint(__ldg(lut + (c < 31u ? c : 31u)));
                    }
                }
            }
            if (!kCompact) fence_proxy_async_smem();
            __syncthreads();

            if (kCompact) {
                // The sweep's list is one stream of 32-bit words, SPARSE and CONTIGUOUS: a touched cell is
                //   { cell index, plane masks [G = ceil(planes / 32)], the non-zero planes' values in plane order }
                // (a touched cell of a KITTI sweep has 4-5 non-zero planes of 27). A tile's cells are in ASCENDING order;
                // the tile takes its words from the sweep's counter (one atomic per non-empty tile) and records
                // {first word, cells} in the sweep's table, so the read-back copies nothing but finished non-zero values
                // and the host rebuilds the view in one linear walk (host_expand.cu, expand_sorted).
                unsigned *out = compact + (size_t)f * compact_cap;
                unsigned *tab = compact_table + (size_t)f * (2u * (unsigned)tpf + 1u);
                const int G = (planes + 31) >> 5;
                __shared__ unsigned compact_base;
                // (1) words per touched cell, left in the cell's (now dead) winner key; untouched cells hold 0
                for (int d = warp; d < nd; d += kWarps) {
                    const unsigned local = dirty[d];
                    unsigned words = 1u + (unsigned)G;
                    for (int g = 0; g < G; g++) {
                        const int k = g * 32 + lane;
                        const unsigned v = k < planes ? img[local * planes + k] : 0u;
                        words += (unsigned)__popc(__ballot_sync(0xFFFFFFFFu, v != 0u));
                    }
                    if (lane == 0) key[local] = (unsigned long long)words;
                }
                __syncthreads();
                // (2) exclusive prefix over the tile's cells in cell order (one warp), the tile's words from the counter
                if (warp == 0) {
                    unsigned carry = 0;
                    for (int c0 = 0; c0 < T; c0 += 32) {
                        const int c = c0 + lane;
                        const unsigned w = c < T ? (unsigned)key[c] : 0u;
                        unsigned incl = w;
#pragma unroll
                        for (int o = 1; o < 32; o <<= 1) {
                            const unsigned up = __shfl_up_sync(0xFFFFFFFFu, incl, o);
                            if (lane >= o) incl += up;
                        }
                        if (w) key[c] = ((unsigned long long)(carry + incl - w) << 32) | (unsigned long long)w;
                        carry += __shfl_sync(0xFFFFFFFFu, incl, 31);
                    }
                    if (lane == 0) {
                        compact_base = atomicAdd(tab + 2u * (unsigned)tpf, carry);
                        tab[2 * tl] = compact_base;
                        tab[2 * tl + 1] = (unsigned)nd;
                    }
                }
                __syncthreads();
                // (3) one warp per touched cell writes its entry
                for (int d = warp; d < nd; d += kWarps) {
                    const unsigned local = dirty[d];
                    const unsigned long long kk = key[local];
                    const unsigned pos = compact_base + (unsigned)(kk >> 32);
                    MV3D_CHECK(pos + (unsigned)kk <= compact_cap, MV3D_BAD_STREAM);
                    if (pos + (unsigned)kk > compact_cap) continue;  // the stream holds every point's worst case: never
                    if (lane == 0) out[pos] = (unsigned)cell0 + local;
                    unsigned at = pos + 1u + (unsigned)G;
                    for (int g = 0; g < G; g++) {
                        const int k = g * 32 + lane;
                        const unsigned v = k < planes ? img[local * planes + k] : 0u;
                        const unsigned m = __ballot_sync(0xFFFFFFFFu, v != 0u);
                        if (lane == 0) out[pos + 1u + (unsigned)g] = m;
                        if (v != 0u) out[at + (unsigned)__popc(m & ((1u << lane) - 1u))] = v;
                        at += (unsigned)__popc(m);
                    }
                }
            } else if (kStore == kStorePatch) {
                for (int d = warp; d < nd; d += kWarps) {  // one warp per touched cell: its planes are contiguous
                    const unsigned local = dirty[d];
                    for (int k = lane; k < planes; k += 32) dst[local * planes + k] = __uint_as_float(img[local * planes + k]);
                }
            } else if (bulk) {
                if (tid == 0) { bulk_store_hint(dst, img, bytes, l2_policy_evict_first()); bulk_commit(); bulk_wait_read_all(); }
            } else {
                for (int w = tid; w < ncell * planes; w += kTileThreads) dst[w] = __uint_as_float(img[w]);
            }
            if (kOccupancy) {
                const unsigned w_first = (unsigned)cell0 >> 5, w_last = ((unsigned)cell0 + (unsigned)ncell - 1u) >> 5;
                unsigned *out = sink.words + (size_t)f * sink.words_per_frame + w_first;
                for (unsigned w = tid; w <= w_last - w_first; w += kTileThreads) {
                    const unsigned bits = occ_tile[w];
                    if (bits) {
                        // a tile of a multiple of 32 cells owns all its words; otherwise its first and last word
                        // are shared with the neighbouring tiles
                        if ((T & 31) != 0 && (w == 0 || w == w_last - w_first)) atomicOr(out + w, bits);
                        else out[w] = bits;
                        occ_tile[w] = 0u;
                    }
                }
            }
            __syncthreads();
            for (int d = warp; d < nd; d += kWarps) {
                const unsigned local = dirty[d];
                for (int k = lane; k < planes; k += 32) img[local * planes + k] = 0u;
                if (lane == 0) key[local] = 0ull;
            }
            if (tid == 0) ndirty = 0;
            if (!kCompact) fence_proxy_async_smem();
            __syncthreads();
        }
        rec0 = rec1; n0 = n1; q0 = q1;
        rec1 = rec2; n1 = n2;
    }
    if (!kCompact && tid == 0) bulk_wait_read_all();  // shared memory must outlive the last bulk read
}

template <bool kPy>
__global__ void __launch_bounds__(kTileThreads, MV3D_TILE_MIN_BLOCKS)
tile_build_kernel(const uint4 *__restrict__ binned, const unsigned *__restrict__ tile_off,
                  const __grid_constant__ FrameOffsets offs,
                  const __grid_constant__ ViewParams p, const float *__restrict__ lut, float *__restrict__ top,
                  int total_tiles)
{
    tile_body<false, false, kPy>(binned, tile_off, offs, p, lut, top, total_tiles, nullptr, 0u);
}

// Same, and the occupancy bits / flag of every frame for the empty-anchor filter (see tile_body).
template <bool kPy>
__global__ void __launch_bounds__(kTileThreads, MV3D_TILE_MIN_BLOCKS)
tile_build_occupancy_kernel(const uint4 *__restrict__ binned, const unsigned *__restrict__ tile_off,
                  const __grid_constant__ FrameOffsets offs,
                            const __grid_constant__ ViewParams p, const float *__restrict__ lut,
                            float *__restrict__ top, int total_tiles, OccupancySink sink)
{
    tile_body<false, true, kPy>(binned, tile_off, offs, p, lut, top, total_tiles, nullptr, 0u, sink);
}

// Same reduction for an output the copy engine has zero-filled: only the touched cells are written (PATCH strategy).
template <bool kPy>
__global__ void __launch_bounds__(kTileThreads, MV3D_TILE_MIN_BLOCKS)
tile_patch_kernel(const uint4 *__restrict__ binned, const unsigned *__restrict__ tile_off,
                  const __grid_constant__ FrameOffsets offs,
                  const __grid_constant__ ViewParams p, const float *__restrict__ lut, float *__restrict__ top,
                  int total_tiles)
{
    tile_body<false, false, kPy, kStorePatch>(binned, tile_off, offs, p, lut, top, total_tiles, nullptr, 0u);
}

// Same reduction, compact output (see tile_body).
template <bool kPy>
__global__ void __launch_bounds__(kTileThreads, MV3D_TILE_MIN_BLOCKS)
tile_compact_kernel(const uint4 *__restrict__ binned, const unsigned *__restrict__ tile_off,
                  const __grid_constant__ FrameOffsets offs,
                    const __grid_constant__ ViewParams p, const float *__restrict__ lut, int total_tiles,
                    unsigned *__restrict__ compact, unsigned compact_cap, unsigned *__restrict__ compact_table)
{
    tile_body<true, false, kPy>(binned, tile_off, offs, p, lut, nullptr, total_tiles, compact, compact_cap, OccupancySink(),
                                compact_table);
}


// ------------------------------------------------------------------------------------------------
// Front view: write the whole [R][C][F] image from the per-pixel winners, and clear the winners.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kPointThreads)
front_finalize_kernel(const float4 *__restrict__ pts, const __grid_constant__ FrameOffsets offs,
                      const __grid_constant__ ViewParams p, unsigned *__restrict__ fv_winner, float *__restrict__ front)
{
    pdl_trigger();
    pdl_wait();  // the winners are the output of the pass in front of this launch
    front_finalize_body(pts, offs, p, fv_winner, front, blockIdx.y, blockIdx.x, gridDim.x);
}

// ------------------------------------------------------------------------------------------------
// Python-twin front view (src/data.py:113-180 lidar_to_front_fast): per pixel, the MEAN of the channel values of
// its points, accumulated by the reference as a float32 running sum in input order.
//   link   : every point that lands in the image is pushed on its pixel's list (atomicExch on the list head)
//   reduce : one thread per pixel walks its list, puts the (few) point indices in increasing order, replays them
//            -- values recomputed from the points, which sit in L2 -- and writes the whole image, zeros included.
// No histogram, scan or scatter: a 1500 x 100 image would need a 150 k-entry scan per sweep for ~50 k points.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kPointThreads)
pyfront_link_kernel(const float4 *__restrict__ pts, FrameOffsets offs, ViewParams p, PyFrontParams fp,
                    unsigned *__restrict__ head, unsigned *__restrict__ next)
{
    const int f = blockIdx.y;
    const int base = offs.off[f];
    const int n = offs.off[f + 1] - base;
    unsigned *heads = head + (size_t)f * fp.pixels;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 pt = load_point(pts + base + i);
        if (!py_crop(pt, p)) continue;
        const int pix = py_front_pixel(pt, fp);
        if (pix >= 0) next[base + i] = atomicExch(heads + pix, (unsigned)i + 1u);  // 0 ends a list
    }
}

__global__ void __launch_bounds__(kPointThreads)
pyfront_reduce_kernel(const float4 *__restrict__ pts, FrameOffsets offs, PyFrontParams fp, unsigned *__restrict__ head,
                      const unsigned *__restrict__ next, float *__restrict__ front)
{
    constexpr int kSmall = 16;
    const int f = blockIdx.y;
    const int base = offs.off[f];
    unsigned *heads = head + (size_t)f * fp.pixels;
    float *out = front + (size_t)f * fp.pixels * 3;
    for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < fp.pixels; pix += gridDim.x * blockDim.x) {
        const unsigned h = heads[pix];
        float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f;
        if (h) {
            heads[pix] = 0u;  // self-cleaning
            unsigned idx[kSmall];
            int n = 0;
            for (unsigned e = h; e; e = next[base + (int)e - 1]) {
                if (n < kSmall) idx[n] = e - 1u;
                n++;
            }
            auto add = [&](unsigned i) {  // `front[c, r, :] += x[2:]` (:172-177): float32, one point at a time
                const float3 v = py_front_values(pts[base + (int)i], fp);
                s0 = __fadd_rn(s0, v.x); s1 = __fadd_rn(s1, v.y); s2 = __fadd_rn(s2, v.z);
            };
            if (n <= kSmall) {
                for (int a = 1; a < n; a++) {  // insertion sort by point index = input order
                    const unsigned v = idx[a];
                    int b = a - 1;
                    while (b >= 0 && idx[b] > v) { idx[b + 1] = idx[b]; b--; }
                    idx[b + 1] = v;
                }
                for (int a = 0; a < n; a++) add(idx[a]);
            } else {
                long long last = -1;  // crowded pixel: repeatedly take the smallest index above the last one
                for (int k = 0; k < n; k++) {
                    unsigned best = 0xFFFFFFFFu;
                    for (unsigned e = h; e; e = next[base + (int)e - 1])
                        if ((long long)(e - 1u) > last && e - 1u < best) best = e - 1u;
                    add(best);
                    last = best;
                }
            }
            const float c = (float)n;  // `front /= weight_mask` (:178)
            s0 = __fdiv_rn(s0, c); s1 = __fdiv_rn(s1, c); s2 = __fdiv_rn(s2, c);
        }
        out[(size_t)pix * 3 + 0] = s0;
        out[(size_t)pix * 3 + 1] = s1;
        out[(size_t)pix * 3 + 2] = s2;
    }
}

// ------------------------------------------------------------------------------------------------
// ATOMIC strategy
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) fill_zero_kernel(float *__restrict__ dst, size_t count)
{
    pdl_trigger();  // the atomic pass behind it may load and classify its first points while this fill drains
    // head (to 16-byte alignment), vector body, tail
    const uintptr_t addr = reinterpret_cast<uintptr_t>(dst);
    size_t head = ((16 - (addr & 15)) & 15) / 4;
    if (head > count) head = count;
    const size_t body = (count - head) / 4;
    float4 *v = reinterpret_cast<float4 *>(dst + head);
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t nthreads = (size_t)gridDim.x * blockDim.x;
    for (size_t i = tid; i < body; i += nthreads) store_zero16(v + i);
    if (tid < head) dst[tid] = 0.0f;
    const size_t tail0 = head + body * 4;
    if (tid < count - tail0) dst[tail0 + tid] = 0.0f;
}

#ifndef MV3D_ATOMIC_ITEMS
#define MV3D_ATOMIC_ITEMS 2
#endif
constexpr int kAtomicItems = MV3D_ATOMIC_ITEMS;  // points per thread and round of the two atomic passes

// Pass 1: reduce straight into the (zeroed) output. Height slots hold kRawTag | h bits until finalised,
// the density slot holds the raw count, cellkey holds the (h, ~index) winner key.
__global__ void __launch_bounds__(kPointThreads)
atomic_scatter_kernel(const float4 *__restrict__ pts, FrameOffsets offs, ViewParams p,
                      float *__restrict__ top, unsigned long long *__restrict__ cellkey,
                      unsigned *__restrict__ fv_winner)
{
    pdl_trigger();
    const int f = blockIdx.y;
    const int base = offs.off[f];
    const int n = offs.off[f + 1] - base;
    unsigned *out = reinterpret_cast<unsigned *>(top + (size_t)f * p.top_frame_floats);
    unsigned long long *keys = cellkey + (size_t)f * ((size_t)p.xn * p.yn);
    unsigned *winners = fv_winner + (size_t)f * p.front_pixels;
    // kAtomicItems points per thread and round: all loads are issued before the first atomic
    const int stride = gridDim.x * blockDim.x;
    bool waited = false;
    for (int i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += kAtomicItems * stride) {
        float4 pt[kAtomicItems];
        CellHit c[kAtomicItems];
#pragma unroll
        for (int k = 0; k < kAtomicItems; k++)
            if (i0 + k * stride < n) pt[k] = load_point(pts + base + i0 + k * stride);
#pragma unroll
        for (int k = 0; k < kAtomicItems; k++) {
            c[k].keep = false;
            if (i0 + k * stride < n) c[k] = classify(pt[k], p);
        }
        if (!waited) {  // the points are the caller's; the zero-fill in front of this launch must be complete from here on
            pdl_wait();
            waited = true;
        }
#pragma unroll
        for (int k = 0; k < kAtomicItems; k++) {
            const int i = i0 + k * stride;
            if (!c[k].keep) continue;
            if (p.want_top) {
                MV3D_CHECK((long long)c[k].cell < (long long)p.xn * p.yn, MV3D_BAD_CELL);
                MV3D_CHECK(c[k].slice >= 0 && c[k].slice < p.zn, MV3D_BAD_SLICE);
                unsigned *cellw = out + (size_t)c[k].cell * p.planes;
                atomicAdd(cellw + p.zn + 1, 1u);
                if (c[k].h > 0.0f) {
                    atomicMax(cellw + c[k].slice, __float_as_uint(c[k].h) | kRawTag);
                    atomicMax(keys + c[k].cell, cell_key(c[k].h, (uint32_t)i));
                }
            }
            if (p.want_front) {
                const int pix = front_pixel(pt[k], p);
                MV3D_CHECK(pix < p.front_pixels, MV3D_BAD_PIXEL);
                if (pix >= 0) atomicMax(winners + pix, (unsigned)i + 1u);
            }
        }
    }
    if (!waited) pdl_wait();
}

// Pass 2: every kept point revisits its cell. The slot winner (tagged value == its own) converts the height;
// the key winner writes the intensity; anyone who still sees a raw count converts it.
// Raw counts (< 2^30) and density floats (>= 0.2f = 0x3E4CCCCD) cannot be confused, and all writers of one
// density slot write the same value, so the race between them is benign.
__global__ void __launch_bounds__(kPointThreads)
atomic_finalize_kernel(const float4 *__restrict__ pts, FrameOffsets offs, ViewParams p, const float *__restrict__ lut,
                       float *__restrict__ top, unsigned long long *__restrict__ cellkey)
{
    pdl_trigger();
    const int f = blockIdx.y;
    const int base = offs.off[f];
    const int n = offs.off[f + 1] - base;
    unsigned *out = reinterpret_cast<unsigned *>(top + (size_t)f * p.top_frame_floats);
    unsigned long long *keys = cellkey + (size_t)f * ((size_t)p.xn * p.yn);
    const int stride = gridDim.x * blockDim.x;
    bool waited = false;
    for (int i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += kAtomicItems * stride) {
        float4 pt[kAtomicItems];
#pragma unroll
        for (int k = 0; k < kAtomicItems; k++)
            if (i0 + k * stride < n) pt[k] = load_point(pts + base + i0 + k * stride);
        // every point's cell words are requested before any is examined
        CellHit c[kAtomicItems];
        unsigned cnt[kAtomicItems], slot[kAtomicItems];
        unsigned long long key[kAtomicItems];
#pragma unroll
        for (int k = 0; k < kAtomicItems; k++) {
            c[k].keep = false;
            if (i0 + k * stride < n) c[k] = classify(pt[k], p);
        }
        if (!waited) {  // the scatter pass in front of this launch must be complete before its results are read
            pdl_wait();
            waited = true;
        }
#pragma unroll
        for (int k = 0; k < kAtomicItems; k++) {
            cnt[k] = slot[k] = 0u;
            key[k] = 0ull;
            if (!c[k].keep) continue;
            unsigned *cellw = out + (size_t)c[k].cell * p.planes;
            cnt[k] = *reinterpret_cast<volatile unsigned *>(cellw + p.zn + 1);
            if (c[k].h > 0.0f) {
                slot[k] = *reinterpret_cast<volatile unsigned *>(cellw + c[k].slice);
                key[k] = *reinterpret_cast<volatile unsigned long long *>(keys + c[k].cell);
            }
        }
#pragma unroll
        for (int k = 0; k < kAtomicItems; k++) {
            if (!c[k].keep) continue;
            const int i = i0 + k * stride;
            unsigned *cellw = out + (size_t)c[k].cell * p.planes;
            if (cnt[k] < 0x3E000000u) cellw[p.zn + 1] = __float_as_uint(__ldg(lut + (cnt[k] < 31u ? cnt[k] : 31u)));
            if (c[k].h > 0.0f) {
                if (slot[k] == (__float_as_uint(c[k].h) | kRawTag))
                    cellw[c[k].slice] = __float_as_uint(slice_value(c[k].h, p.z_div, c[k].slice));
                if (key[k] == cell_key(c[k].h, (uint32_t)i)) cellw[p.zn] = __float_as_uint(pt[k].w);
            }
        }
    }
    if (!waited) pdl_wait();
}

// ------------------------------------------------------------------------------------------------
// Debug / test hook: the per-point decisions, so tests can compare them one by one with the oracle.
//   cell[i]      output cell of point i, or -1 if the gate rejects it
//   pix[i]       front-view pixel through the production path (float fast path + exact fallback), or -1
//   pix_exact[i] the same through the all-double evaluation
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kPointThreads)
point_keys_kernel(const float4 *__restrict__ pts, int n, ViewParams p, int *__restrict__ cell,
                  int *__restrict__ pix, int *__restrict__ pix_exact)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 pt = load_point(pts + i);
        const CellHit c = classify(pt, p);
        cell[i] = c.keep ? (int)c.cell : -1;
        if (p.want_front) {
            pix[i] = c.keep ? front_pixel(pt, p) : -1;
            pix_exact[i] = c.keep ? front_pixel_exact(pt, p) : -1;
        }
    }
}

} // namespace mv3d
namespace mv3d {

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
// Kernel launch with the programmatic-serialization attribute: the kernel may begin (up to its pdl_wait()) while the
// previous kernel in the stream drains. MV3D_PDL=0 falls back to plain stream order (A/B measurements).
static bool pdl_enabled()
{
    static const bool on = [] { const char *e = getenv("MV3D_PDL"); return !e || atoi(e) != 0; }();
    return on;
}
template <typename... Params, typename... Args>
static void launch_after(void (*kernel)(Params...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args &&...args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, static_cast<Params>(args)...);
}

static inline int point_blocks(int max_points, int items = 1)
{
    int b = (max_points + kPointThreads * items - 1) / (kPointThreads * items);
    return b < 1 ? 1 : (b > 4096 ? 4096 : b);
}

// CTAs per sweep for the grid-stride point passes of a batch: enough to cover the longest sweep, but no more than
// MV3D_POINT_WAVES (default 4) waves of resident CTAs over the whole batch. A CTA that lives for one round of 512
// points spends a quarter of its instructions on its prologue (ncu, profiles/r2b_*): with 64 sweeps per launch each
// CTA now loops ~6 times over its sweep instead.
static int sm_count_cached()
{
    static const int sms = [] {
        int dev = 0, n = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n < 1) n = 148;
        return n;
    }();
    return sms;
}
static inline int point_blocks_batched(int max_points, int items, int batch, int ctas_per_sm)
{
    static const int waves = [] { const char *e = getenv("MV3D_POINT_WAVES"); const int w = e ? atoi(e) : 4; return w < 1 ? 1 << 20 : w; }();
    const int cover = point_blocks(max_points, items);
    long long cap = ((long long)sm_count_cached() * ctas_per_sm * waves + batch - 1) / (batch > 0 ? batch : 1);
    if (cap < 1) cap = 1;
    return cover < cap ? cover : (int)cap;
}

size_t tile_smem_bytes(const ViewParams &p)
{
    const size_t img = ((size_t)p.tile_cells * p.planes * 4 + 15) & ~(size_t)15;
    return img + (size_t)p.tile_cells * 8 + (size_t)p.tile_cells * 4 + ((size_t)p.tile_cells / 32 + 2) * 4;
}

// The opt-in is a property of the (device, kernel) pair, shared by every context of the process: only ever raise it,
// or a context configured for a small tile would lower it under another one's cached configuration.
template <bool kPy>
static cudaError_t tile_kernels_opt_in(int smem)
{
    const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
    cudaError_t e;
    if ((e = cudaFuncSetAttribute(tile_build_kernel<kPy>, attr, smem)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(tile_build_occupancy_kernel<kPy>, attr, smem)) != cudaSuccess) return e;
    if ((e = cudaFuncSetAttribute(tile_patch_kernel<kPy>, attr, smem)) != cudaSuccess) return e;
    return cudaFuncSetAttribute(tile_compact_kernel<kPy>, attr, smem);
}

cudaError_t tile_kernels_configure(const ViewParams &p, int *blocks_per_sm)
{
    constexpr int kMaxDevices = 64;
    static std::mutex mu;
    static int opted[kMaxDevices][2] = {};
    const int smem = (int)tile_smem_bytes(p);
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return e;
    {
        std::lock_guard<std::mutex> lock(mu);
        int &have = opted[device < kMaxDevices ? device : kMaxDevices - 1][p.py ? 1 : 0];
        if (smem > have || device >= kMaxDevices) {
            e = p.py ? tile_kernels_opt_in<true>(smem) : tile_kernels_opt_in<false>(smem);
            if (e != cudaSuccess) return e;
            have = smem;
        }
    }
    return p.py ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, tile_build_kernel<true>, kTileThreads, smem)
                : cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, tile_build_kernel<false>, kTileThreads, smem);
}

int launch_classify(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, int batch, int max_points,
                    uint2 *keys, unsigned *tile_count, unsigned *fv_winner, cudaStream_t s)
{
    dim3 grid(point_blocks_batched(max_points, kPointItems, batch, MV3D_CLASSIFY_MIN_BLOCKS), batch);
    auto kernel = p.py ? classify_kernel<true> : classify_kernel<false>;
    kernel<<<grid, kPointThreads, 0, s>>>(pts, offs, p, keys, tile_count, fv_winner);
    return 1;
}

int launch_tile_scan(unsigned *tile_count, unsigned *tile_off, int tiles_per_frame, int batch, cudaStream_t s)
{
    if (tiles_per_frame > 16 * kScanThreads)
        launch_after(tile_scan_kernel<8>, dim3(batch), dim3(kScanThreads), 0, s, tile_count, tile_off, tiles_per_frame);
    else
        launch_after(tile_scan_kernel<2>, dim3(batch), dim3(kScanThreads), 0, s, tile_count, tile_off, tiles_per_frame);
    return 1;
}

int launch_bin(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, int batch, int max_points,
               const uint2 *keys, const unsigned *tile_off, uint4 *binned, cudaStream_t s, unsigned *fv_winner,
               float *front)
{
    const int front_blocks = (front && fv_winner && p.front_pixels > 0) ? point_blocks(p.front_pixels) : 0;
    dim3 grid(point_blocks_batched(max_points, kPointItems, batch, 8) + front_blocks, batch);
    auto kernel = p.py ? bin_kernel<true> : bin_kernel<false>;
    launch_after(kernel, grid, dim3(kPointThreads), 0, s, pts, offs, p, keys, tile_off, binned, fv_winner, front, front_blocks);
    return 1;
}

int launch_tile_build(const uint4 *binned, const unsigned *tile_off, const FrameOffsets &offs, const ViewParams &p,
                      const float *lut, float *top, int batch, int grid_blocks, cudaStream_t s, const OccupancySink *sink)
{
    const int total = batch * p.tiles_per_frame;
    const int grid = grid_blocks < total ? grid_blocks : total;
    if (sink && sink->words) {
        auto kernel = p.py ? tile_build_occupancy_kernel<true> : tile_build_occupancy_kernel<false>;
        launch_after(kernel, dim3(grid), dim3(kTileThreads), tile_smem_bytes(p), s, binned, tile_off, offs, p, lut, top, total, *sink);
    } else {
        auto kernel = p.py ? tile_build_kernel<true> : tile_build_kernel<false>;
        launch_after(kernel, dim3(grid), dim3(kTileThreads), tile_smem_bytes(p), s, binned, tile_off, offs, p, lut, top, total);
    }
    return 1;
}

int launch_tile_patch(const uint4 *binned, const unsigned *tile_off, const FrameOffsets &offs, const ViewParams &p,
                      const float *lut, float *top, int batch, int grid_blocks, cudaStream_t s)
{
    const int total = batch * p.tiles_per_frame;
    const int grid = grid_blocks < total ? grid_blocks : total;
    auto kernel = p.py ? tile_patch_kernel<true> : tile_patch_kernel<false>;
    kernel<<<grid, kTileThreads, tile_smem_bytes(p), s>>>(binned, tile_off, offs, p, lut, top, total);
    return 1;
}

int launch_tile_compact(const uint4 *binned, const unsigned *tile_off, const FrameOffsets &offs, const ViewParams &p,
                        const float *lut, int batch, int grid_blocks, unsigned *compact, unsigned compact_cap,
                        unsigned *compact_table, cudaStream_t s)
{
    const int total = batch * p.tiles_per_frame;
    const int grid = grid_blocks < total ? grid_blocks : total;
    auto kernel = p.py ? tile_compact_kernel<true> : tile_compact_kernel<false>;
    launch_after(kernel, dim3(grid), dim3(kTileThreads), tile_smem_bytes(p), s, binned, tile_off, offs, p, lut, total, compact, compact_cap,
                 compact_table);
    return 1;
}

int launch_front_finalize(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, int batch,
                          unsigned *fv_winner, float *front, cudaStream_t s)
{
    dim3 grid(point_blocks(p.front_pixels), batch);
    launch_after(front_finalize_kernel, grid, dim3(kPointThreads), 0, s, pts, offs, p, fv_winner, front);
    return 1;
}

int launch_pyfront_link(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, const PyFrontParams &fp,
                        int batch, int max_points, unsigned *head, unsigned *next, cudaStream_t s)
{
    dim3 grid(point_blocks(max_points), batch);
    pyfront_link_kernel<<<grid, kPointThreads, 0, s>>>(pts, offs, p, fp, head, next);
    return 1;
}

int launch_pyfront_reduce(const float4 *pts, const FrameOffsets &offs, const PyFrontParams &fp, int batch, unsigned *head,
                          const unsigned *next, float *front, cudaStream_t s)
{
    dim3 grid(point_blocks(fp.pixels), batch);
    pyfront_reduce_kernel<<<grid, kPointThreads, 0, s>>>(pts, offs, fp, head, next, front);
    return 1;
}

int launch_point_keys(const float4 *pts, int n, const ViewParams &p, int *cell, int *pix, int *pix_exact, cudaStream_t s)
{
    if (n <= 0) return 0;
    point_keys_kernel<<<point_blocks(n), kPointThreads, 0, s>>>(pts, n, p, cell, pix, pix_exact);
    return 1;
}

int launch_fill_zero(float *dst, size_t count, int sm_count, cudaStream_t s)
{
    if (count == 0) return 0;
    size_t want = (count / 4 + 255) / 256;
    const size_t cap = (size_t)sm_count * 16;
    int blocks = (int)(want < 1 ? 1 : (want > cap ? cap : want));
    fill_zero_kernel<<<blocks, 256, 0, s>>>(dst, count);
    return 1;
}

int launch_atomic_scatter(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, int batch, int max_points,
                          float *top, unsigned long long *cellkey, unsigned *fv_winner, cudaStream_t s)
{
    dim3 grid(point_blocks(max_points), batch);
    launch_after(atomic_scatter_kernel, grid, dim3(kPointThreads), 0, s, pts, offs, p, top, cellkey, fv_winner);
    return 1;
}

int launch_atomic_finalize(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, const float *lut,
                           int batch, int max_points, float *top, unsigned long long *cellkey, cudaStream_t s)
{
    dim3 grid(point_blocks(max_points), batch);
    launch_after(atomic_finalize_kernel, grid, dim3(kPointThreads), 0, s, pts, offs, p, lut, top, cellkey);
    return 1;
}

MV3D_SELF_CHECK_READER(selfcheck_encode)
__global__ void selfcheck_trip_kernel(int bit) { MV3D_CHECK(bit < 0, MV3D_BAD_ON_PURPOSE); }
void selfcheck_trip(cudaStream_t s) { selfcheck_trip_kernel<<<1, 1, 0, s>>>(1); }

} // namespace mv3d
