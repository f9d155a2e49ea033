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
//  PATCH           fill_tma (side stream)  ||  classify -> scan -> bin ; then warp_patch per chunk of sweeps
//      The zero-fill of the output needs no point data and almost no SM resources when it is done by bulk
//      async (TMA) stores from one all-zero shared-memory buffer, so it runs beside the compute/latency-bound
//      point passes instead of after them. Points are binned by small tiles; one WARP then owns a tile:
//      it reduces the tile's records in a private shared-memory image and overwrites only the 32-byte
//      sectors of the (already zero) output that hold a touched cell.
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
// How the tile kernel writes a finished tile (ViewParams::store_mode; MV3D_TILE_STORE=tma|tma_hint|tma_split|lsu|lsu_cs).
// Default: bulk stores with an L2 evict-first policy -- the output is written once and never read back here, so it
// should not push the bin records (read by this same kernel) out of L2. Measured in profiles/strategy_experiments_r1.md.
// kStorePatch: the output has been zero-filled already (copy engine); write the touched cells only.
enum TileStore { kStoreTma = 0, kStoreTmaHint = 1, kStoreTmaSplit = 2, kStoreLsu = 3, kStoreLsuStream = 4, kStoreTmaPaced = 5, kStorePatch = 6 };
__device__ __forceinline__ void tile_bulk_store(void *gdst, const void *ssrc, uint32_t bytes, int mode)
{
    if (mode == kStoreTmaHint) {
        bulk_store_hint(gdst, ssrc, bytes, l2_policy_evict_first());
    } else if (mode == kStoreTmaPaced) {  // four bulk operations, each waited for: at most a quarter tile in flight
        const uint32_t quarter = ((bytes >> 2) + 15u) & ~15u;
        const uint64_t pol = l2_policy_evict_first();
        for (uint32_t off = 0; off < bytes; off += quarter) {
            bulk_store_hint(static_cast<char *>(gdst) + off, static_cast<const char *>(ssrc) + off,
                            min(quarter, bytes - off), pol);
            bulk_commit();
            bulk_wait_read_all();
        }
    } else if (mode == kStoreTmaSplit) {  // four bulk operations per tile instead of one
        const uint32_t quarter = ((bytes >> 2) + 15u) & ~15u;
        for (uint32_t off = 0; off < bytes; off += quarter)
            bulk_store(static_cast<char *>(gdst) + off, static_cast<const char *>(ssrc) + off, min(quarter, bytes - off));
    } else {
        bulk_store(gdst, ssrc, bytes);
    }
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
                        if (pix >= 0) atomicMax(winners + pix, (unsigned)idx[k] + 1u);
                    }
                }
            }
            if (cell[k] != kInvalidKey) tile[k] = tile_of(cell[k], p);
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
// TILE strategy, pass 3: move each kept point's record into its tile's bin.
//   record = { cell-in-tile | slice << 16,  h bits,  point index,  intensity bits }   (16 bytes)
// Order inside a bin is arbitrary; every order-dependent decision uses the point index carried along.
// ------------------------------------------------------------------------------------------------
template <bool kPy>
__global__ void __launch_bounds__(kPointThreads)
bin_kernel(const float4 *__restrict__ pts, const __grid_constant__ FrameOffsets offs,
           const __grid_constant__ ViewParams p, const uint2 *__restrict__ keys,
           const unsigned *__restrict__ tile_off, uint4 *__restrict__ binned)
{
    const int f = blockIdx.y;
    const int base = offs.off[f];
    const int n = offs.off[f + 1] - base;
    const unsigned *off = tile_off + (size_t)f * (p.tiles_per_frame + 1);
    const int stride = gridDim.x * blockDim.x;
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
//   compact[f] = { cell index [cap] , values [cap][planes] }   (structure of arrays)
// A tile with n records touches at most n cells, so it owns the slots [beg, beg + n) of its own bin range
// (tile_off): the first nd hold its touched cells (any order), the rest are marked kNoCell. The list is therefore
// ordered by tile with no atomics and no scan, and tile_off tells the reader where each tile's slots are. This is
// what the host-buffer path reads back over PCIe instead of the 95 %-zero dense tensor.
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
template <bool kCompact, bool kOccupancy, bool kPy, int kStore = kStoreTmaHint,
          int kInFlight = (kCompact || kOccupancy || kPy) ? 1 : kRecordsInFlight>
__device__ __forceinline__ void
tile_body(const uint4 *__restrict__ binned, const unsigned *__restrict__ tile_off, const FrameOffsets &offs,
          const ViewParams &p, const float *__restrict__ lut, float *__restrict__ top, int total_tiles,
          unsigned *__restrict__ compact, unsigned compact_cap, OccupancySink sink = OccupancySink())
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
            // nothing to report for an empty tile
        } else if (n0 == 0) {
            // Empty tile: stream the all-zero image; nothing is modified, so no barrier.
            if (kStore == kStorePatch) {
                // the output is zero already
            } else if (bulk && kStore == kStoreLsu) {
                float4 *d4 = reinterpret_cast<float4 *>(dst);
                for (int w = tid; w < (int)(bytes >> 4); w += kTileThreads) d4[w] = make_float4(0.f, 0.f, 0.f, 0.f);
            } else if (bulk && kStore == kStoreLsuStream) {
                float4 *d4 = reinterpret_cast<float4 *>(dst);
                for (int w = tid; w < (int)(bytes >> 4); w += kTileThreads) store_zero16(d4 + w);
            } else if (bulk) {
                if (tid == 0) { tile_bulk_store(dst, img, bytes, kStore); bulk_commit(); }
            } else {
                for (int w = tid; w < ncell * planes; w += kTileThreads) dst[w] = 0.0f;
            }
        } else {
            // Earlier (empty-tile) bulk stores may still be reading the image: wait before touching it.
            if (!kCompact && kStore < kStoreLsu) {
                if (tid == 0) bulk_wait_read_all();
                __syncthreads();
            }

            // accumulate: the first kTileThreads records are in registers already (prefetched); the rest of a heavy
            // tile (tiles next to the sensor hold hundreds to thousands of records) are fetched four per thread at
            // a time, so that their L2 latency overlaps
            auto accumulate = [&](const uint4 &q) {
                const unsigned local = q.x & 0xFFFFu, Z = q.x >> 16;
                unsigned *cellw = img + local * planes;
                if (atomicAdd(cellw + zn + 1, 1u) == 0u) dirty[atomicAdd(&ndirty, 1u)] = local;
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
                // the tile's slots are those of its records: touched cells first, one warp per cell
                unsigned *out_idx = compact + (size_t)f * compact_cap * ((unsigned)planes + 1u);
                unsigned *out_val = out_idx + compact_cap;
                const unsigned beg = (unsigned)(rec0 - (binned + (size_t)offs.off[f]));
                for (int d = warp; d < nd; d += kWarps) {
                    const unsigned slot = beg + (unsigned)d;
                    if (slot >= compact_cap) continue;  // cap >= points of the sweep: cannot happen, never overrun
                    const unsigned local = dirty[d];
                    if (lane == 0) out_idx[slot] = (unsigned)cell0 + local;
                    for (int k = lane; k < planes; k += 32) out_val[(size_t)slot * planes + k] = img[local * planes + k];
                }
                for (int d = nd + tid; d < n0; d += kTileThreads)
                    if (beg + (unsigned)d < compact_cap) out_idx[beg + (unsigned)d] = kNoCell;
            } else if (kStore == kStorePatch) {
                for (int d = warp; d < nd; d += kWarps) {  // one warp per touched cell: its planes are contiguous
                    const unsigned local = dirty[d];
                    for (int k = lane; k < planes; k += 32) dst[local * planes + k] = __uint_as_float(img[local * planes + k]);
                }
            } else if (bulk && kStore >= kStoreLsu) {
                const uint4 *s4 = reinterpret_cast<const uint4 *>(img);
                uint4 *d4 = reinterpret_cast<uint4 *>(dst);
                if (kStore == kStoreLsu) {
                    for (int w = tid; w < (int)(bytes >> 4); w += kTileThreads) d4[w] = s4[w];
                } else {
                    for (int w = tid; w < (int)(bytes >> 4); w += kTileThreads) __stcs(d4 + w, s4[w]);
                }
            } else if (bulk) {
                if (tid == 0) { tile_bulk_store(dst, img, bytes, kStore); bulk_commit(); bulk_wait_read_all(); }
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

template <bool kPy, int kStore = kStoreTmaHint>
__global__ void __launch_bounds__(kTileThreads, MV3D_TILE_MIN_BLOCKS)
tile_build_kernel(const uint4 *__restrict__ binned, const unsigned *__restrict__ tile_off,
                  const __grid_constant__ FrameOffsets offs,
                  const __grid_constant__ ViewParams p, const float *__restrict__ lut, float *__restrict__ top,
                  int total_tiles)
{
    tile_body<false, false, kPy, kStore>(binned, tile_off, offs, p, lut, top, total_tiles, nullptr, 0u);
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
                    unsigned *__restrict__ compact, unsigned compact_cap)
{
    tile_body<true, false, kPy>(binned, tile_off, offs, p, lut, nullptr, total_tiles, compact, compact_cap);
}


// ------------------------------------------------------------------------------------------------
// TILE strategy, pass 4, pipelined form: ONE CTA per SM, G tile images in shared memory, G builder groups and one
// store warp (MV3D_TILE_PIPE).
//
// Why: the bulk-store stream runs fastest when each SM keeps only ~9 KB of it in flight (6.8 TB/s, against 6.2-6.4
// with whole 36 KB tiles of five independent CTAs in flight: profiles/tma_store_bench_r1.txt). Pacing the store inside
// tile_build_kernel stalls the CTA that should be building the next tile, so here the two are decoupled: builder
// group g owns image g and reduces tile (k*G + g)*grid + cta into it; the store warp walks the images in that same
// order, writes each finished tile in chunks of pipe_chunk bytes with at most pipe_pending chunks unread, and hands
// the image back. full[g] / empty[g] are mbarriers; a builder re-zeroes the cells its previous tile touched once
// empty[g] says the store has read them. Same arithmetic, same order-preserving atomics as tile_body.
// ------------------------------------------------------------------------------------------------
constexpr int kPipeMaxGroups = 5;  // hardware barriers: 0 = whole CTA, then per group: its own and "empty"
__device__ __forceinline__ void named_sync(int id, int threads)
{
    asm volatile("bar.sync %0, %1;" :: "r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void named_arrive(int id, int threads)
{
    asm volatile("bar.arrive %0, %1;" :: "r"(id), "r"(threads) : "memory");
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_test(uint64_t *bar, unsigned parity)
{
    unsigned ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(smem_addr(bar)), "r"(parity) : "memory");
    return ok != 0u;
}
__device__ __forceinline__ void bulk_wait_read_pending(int n)
{
    switch (n) {
    case 0: asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); break;
    case 1: asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); break;
    case 2: asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory"); break;
    case 3: asm volatile("cp.async.bulk.wait_group.read 3;" ::: "memory"); break;
    case 4: asm volatile("cp.async.bulk.wait_group.read 4;" ::: "memory"); break;
    case 5: asm volatile("cp.async.bulk.wait_group.read 5;" ::: "memory"); break;
    case 6: asm volatile("cp.async.bulk.wait_group.read 6;" ::: "memory"); break;
    default: asm volatile("cp.async.bulk.wait_group.read 7;" ::: "memory"); break;
    }
}

struct PipeLayout {
    size_t img_bytes, key_off, dirty_off, occ_off, stride, header;
};
__host__ __device__ inline PipeLayout pipe_layout(int tile_cells, int planes)
{
    PipeLayout l;
    l.img_bytes = ((size_t)tile_cells * planes * 4 + 15) & ~(size_t)15;
    l.key_off = (l.img_bytes + 7) & ~(size_t)7;
    l.dirty_off = l.key_off + (size_t)tile_cells * 8;
    l.occ_off = l.dirty_off + (size_t)tile_cells * 4;
    l.stride = (l.occ_off + ((size_t)tile_cells / 32 + 2) * 4 + 127) & ~(size_t)127;
    l.header = 256;  // density table (32 floats), cnt[5][2], full[5] mbarriers
    return l;
}

template <bool kOccupancy, bool kPy>
__global__ void __launch_bounds__(1024, 1)
tile_pipe_kernel(const uint4 *__restrict__ binned, const unsigned *__restrict__ tile_off,
                 const __grid_constant__ FrameOffsets offs, const __grid_constant__ ViewParams p,
                 const float *__restrict__ lut, float *__restrict__ top, int total_tiles, OccupancySink sink)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int T = p.tile_cells, planes = p.planes, zn = p.zn;
    const int G = p.pipe_groups, gthreads = p.pipe_group_threads;
    const PipeLayout L = pipe_layout(T, planes);
    float *slut = reinterpret_cast<float *>(smem_raw);
    unsigned *cnt = reinterpret_cast<unsigned *>(smem_raw + 128);  // [G][2] touched cells of the current / previous tile
    uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw + 192);  // full[g]: builder group g has finished its tile
    unsigned char *bufs = smem_raw + L.header;
    const int tid = threadIdx.x;
    const int nbuild = G * gthreads;
    const int hand = gthreads + 32;  // a builder group and the store warp meet on the "full" / "empty" barriers

    {   // all images and keys start all-zero
        uint4 *z = reinterpret_cast<uint4 *>(bufs);
        const int n16 = (int)((L.stride * (size_t)G) >> 4);
        for (int w = tid; w < n16; w += blockDim.x) z[w] = make_uint4(0u, 0u, 0u, 0u);
        if (tid < 2 * kPipeMaxGroups) cnt[tid] = 0u;
        if (tid < kDensityLutSize) slut[tid] = __ldg(lut + tid);
        if (tid < G) mbar_init(full + tid, (unsigned)gthreads);
    }
    fence_proxy_async_smem();
    __syncthreads();

    const int grid = gridDim.x, b = blockIdx.x;
    const int tpf = p.tiles_per_frame;
    const long long cells_total = (long long)p.xn * p.yn;
    const int tstep = G * grid;  // total_tiles + 2 * tstep stays far below 2^31 (checked by the launcher)

    if (tid >= nbuild) {
        // ---------------- store warp ----------------
        // Stores whichever image is finished (build times are skewed: tiles near the sensor hold hundreds of records),
        // polling the groups' mbarriers round-robin. The bulk stores of consecutive tiles form one stream of chunks:
        // before a chunk is issued at most pipe_pending older chunks may still be unread, and an image goes back to
        // its builders (the "empty" barrier) as soon as the stream has moved past its last chunk -- so the next
        // tile is found and addressed while the previous one is still being read.
        const int lane = tid - nbuild;
        const uint64_t pol = l2_policy_evict_first();
        const uint32_t chunk = (uint32_t)p.pipe_chunk;
        const int pending = p.pipe_pending;
        int kq[kPipeMaxGroups];  // next round of each group
#pragma unroll
        for (int g = 0; g < kPipeMaxGroups; g++) kq[g] = 0;
        int left = 0;  // tiles of this CTA
        for (int g = 0; g < G; g++) {
            const int first = g * grid + b;
            if (first < total_tiles) left += (total_tiles - 1 - first) / tstep + 1;
        }
        int prev_g = -1;  // image whose store has been issued but which has not been handed back yet
        auto release = [&](int g) {  // all lanes
            __syncwarp();
            named_arrive(1 + kPipeMaxGroups + g, hand);
        };
        while (left > 0) {
            int idle = 1;
#pragma unroll
            for (int g = 0; g < kPipeMaxGroups; g++) {
                if (g >= G) continue;
                const int k = kq[g];
                const int t = (k * G + g) * grid + b;
                if (t >= total_tiles) continue;
                int ready = (lane == 0) ? (int)mbar_test(full + g, (unsigned)(k & 1)) : 0;
                ready = __shfl_sync(0xffffffffu, ready, 0);
                if (!ready) continue;
                idle = 0;
                kq[g] = k + 1;
                left--;
                const int f = t / tpf, tl = t - f * tpf;
                const long long cell0 = (long long)tl * T;
                const int ncell = (int)min((long long)T, cells_total - cell0);
                const uint32_t bytes = (uint32_t)ncell * (uint32_t)planes * 4u;
                float *dst = top + (size_t)f * p.top_frame_floats + (size_t)cell0 * planes;
                const unsigned char *img = bufs + (size_t)g * L.stride;
                const bool bulk = ((reinterpret_cast<uintptr_t>(dst) | (uintptr_t)bytes) & 15u) == 0;
                const bool again = t + tstep < total_tiles;  // the builders will want this image back
                if (bulk) {
                    int j = 0;
                    for (uint32_t off = 0; off < bytes; off += chunk, j++) {
                        if (lane == 0) bulk_wait_read_pending(pending);
                        if (prev_g >= 0 && j == pending) { release(prev_g); prev_g = -1; }
                        if (lane == 0) {
                            bulk_store_hint(reinterpret_cast<char *>(dst) + off, img + off, min(chunk, bytes - off), pol);
                            bulk_commit();
                        }
                    }
                    if (prev_g >= 0) {  // fewer chunks than the stream may leave unread: wait for all but this tile's
                        if (lane == 0) bulk_wait_read_pending(j);
                        release(prev_g);
                    }
                    prev_g = again ? g : -1;
                } else {
                    if (lane == 0) bulk_wait_read_all();
                    if (prev_g >= 0) { release(prev_g); prev_g = -1; }
                    const unsigned *s = reinterpret_cast<const unsigned *>(img);
                    for (int w = lane; w < ncell * planes; w += 32) dst[w] = __uint_as_float(s[w]);
                    if (again) release(g);
                }
            }
            if (idle && prev_g >= 0) {  // nothing to issue: do not sit on the last image
                if (lane == 0) bulk_wait_read_all();
                release(prev_g);
                prev_g = -1;
            }
        }
        if (lane == 0) bulk_wait_read_all();  // shared memory must outlive the last bulk read
        return;
    }

    // ---------------- builder group g ----------------
    const int g = tid / gthreads, gt = tid - g * gthreads;
    const int lane = gt & 31, gwarp = gt >> 5, gwarps = gthreads >> 5;
    const int bar_own = 1 + g, bar_empty = 1 + kPipeMaxGroups + g;
    unsigned char *buf = bufs + (size_t)g * L.stride;
    unsigned *img = reinterpret_cast<unsigned *>(buf);
    unsigned long long *key = reinterpret_cast<unsigned long long *>(buf + L.key_off);
    unsigned *dirty = reinterpret_cast<unsigned *>(buf + L.dirty_off);
    unsigned *occ_tile = reinterpret_cast<unsigned *>(buf + L.occ_off);
    unsigned *mycnt = cnt + 2 * g;

    auto bin_range = [&](int t, const uint4 *&rec, int &n) {
        n = 0;
        rec = binned;
        if (t < total_tiles) {
            const int f = t / tpf, tl = t - f * tpf;
            const unsigned *off = tile_off + (size_t)f * (tpf + 1) + tl;
            const unsigned beg = __ldg(off);
            n = (int)(__ldg(off + 1) - beg);
            rec = binned + (size_t)offs.off[f] + beg;
        }
    };
    const int t_first = g * grid + b;
    const uint4 *rec0, *rec1, *rec2;
    int n0, n1, n2;
    bin_range(t_first, rec0, n0);
    bin_range(t_first + tstep, rec1, n1);
    uint4 q0 = make_uint4(0, 0, 0, 0), q1 = q0;
    if (gt < n0) q0 = __ldg(rec0 + gt);

    int k = 0;
    for (int t = t_first; t < total_tiles; t += tstep, k++) {
        bin_range(t + 2 * tstep, rec2, n2);
        if (gt < n1) q1 = __ldg(rec1 + gt);
        const int f = t / tpf, tl = t - f * tpf;
        const long long cell0 = (long long)tl * T;
        unsigned *cur = mycnt + (k & 1), *prev = mycnt + ((k & 1) ^ 1);

        if (k > 0) named_sync(bar_empty, hand);  // the store warp has read the previous tile of this image
        const int nd_prev = (int)*prev;
        if (n0 > 0 || nd_prev > 0) {
            // re-zero what the previous tile touched
            for (int d = gwarp; d < nd_prev; d += gwarps) {
                const unsigned local = dirty[d];
                for (int kk = lane; kk < planes; kk += 32) img[local * planes + kk] = 0u;
                if (lane == 0) key[local] = 0ull;
            }
            if (gt == 0) *cur = 0u;
            named_sync(bar_own, gthreads);

            // accumulate
            for (int r = gt; r < n0; r += gthreads) {
                const uint4 q = (r == gt) ? q0 : __ldg(rec0 + r);
                const unsigned local = q.x & 0xFFFFu, Z = q.x >> 16;
                unsigned *cellw = img + local * planes;
                if (atomicAdd(cellw + zn + 1, 1u) == 0u) dirty[atomicAdd(cur, 1u)] = local;
                if ((int)q.y > 0) {
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
            }
            named_sync(bar_own, gthreads);

            // finalise the touched cells in place
            const int nd = (int)*cur;
            for (int r = gt; r < n0; r += gthreads) {
                const uint4 q = (r == gt) ? q0 : __ldg(rec0 + r);
                const unsigned local = q.x & 0xFFFFu;
                if ((int)q.y > 0 &&
                    key[local] == (((unsigned long long)q.y << 32) | (unsigned long long)(0xFFFFFFFFu - q.z))) {
                    img[local * planes + zn] = q.w;
                    if (kOccupancy && ((q.w >> 31) != 0u || (q.w & 0x7fffffffu) > 0x7f800000u))
                        atomicOr(sink.flags + f, 1u);
                }
            }
            for (int d = gwarp; d < nd; d += gwarps) {
                unsigned *cellw = img + dirty[d] * planes;
                if (kOccupancy && lane == 0) {
                    const unsigned gc = (unsigned)cell0 + dirty[d];
                    atomicOr(occ_tile + ((gc >> 5) - ((unsigned)cell0 >> 5)), 1u << (gc & 31u));
                }
                for (int kk = lane; kk < planes; kk += 32) {
                    if (kk < zn) {
                        const unsigned raw = cellw[kk];
                        if (raw) cellw[kk] = __float_as_uint(kPy ? __fsub_rn(__uint_as_float(raw), (float)kk)
                                                                  : slice_value(__uint_as_float(raw), p.z_div, kk));
                    } else if (kk == zn + 1) {
                        const unsigned c = cellw[kk];
                        cellw[kk] = __float_as_uint(slut[c < 31u ? c : 31u]);
                    }
                }
            }
            if (kOccupancy) {
                named_sync(bar_own, gthreads);
                const int ncell = (int)min((long long)T, cells_total - cell0);
                const unsigned w_first = (unsigned)cell0 >> 5, w_last = ((unsigned)cell0 + (unsigned)ncell - 1u) >> 5;
                unsigned *out = sink.words + (size_t)f * sink.words_per_frame + w_first;
                for (unsigned w = gt; w <= w_last - w_first; w += gthreads) {
                    const unsigned bits = occ_tile[w];
                    if (bits) {
                        if ((T & 31) != 0 && (w == 0 || w == w_last - w_first)) atomicOr(out + w, bits);
                        else out[w] = bits;
                        occ_tile[w] = 0u;
                    }
                }
            }
        } else if (gt == 0) {
            *cur = 0u;
        }
        fence_proxy_async_smem();
        mbar_arrive(full + g);  // gthreads arrivals complete the phase: the image is ready for the store warp
        rec0 = rec1; n0 = n1; q0 = q1;
        rec1 = rec2; n1 = n2;
    }
}

size_t tile_pipe_smem_bytes(const ViewParams &p)
{
    const PipeLayout l = pipe_layout(p.tile_cells, p.planes);
    return l.header + l.stride * (size_t)p.pipe_groups;
}

// ------------------------------------------------------------------------------------------------
// PATCH strategy, zero-fill: bulk async (TMA) stores from one all-zero shared-memory buffer. One elected thread
// per CTA issues the stores, so the kernel holds one warp and 16 KB of shared memory per CTA and leaves the SMs
// to the point passes running beside it on another stream.
// ------------------------------------------------------------------------------------------------
constexpr int kFillThreads = 32;
constexpr uint32_t kFillChunk = 16384;  // bytes per bulk store
__global__ void __launch_bounds__(kFillThreads) fill_tma_kernel(float *__restrict__ dst, size_t count, int max_pending)
{
    __shared__ __align__(128) float4 zeros[kFillChunk / 16];
    for (int w = threadIdx.x; w < (int)(kFillChunk / 16); w += kFillThreads) zeros[w] = make_float4(0.f, 0.f, 0.f, 0.f);
    fence_proxy_async_smem();
    __syncthreads();
    // head (to 16-byte alignment) and tail by plain stores, body by bulk stores
    const uintptr_t addr = reinterpret_cast<uintptr_t>(dst);
    size_t head = ((16 - (addr & 15)) & 15) / 4;
    if (head > count) head = count;
    const size_t body_bytes = ((count - head) * 4) & ~(size_t)15;
    char *body = reinterpret_cast<char *>(dst + head);
    if (blockIdx.x == 0) {
        const size_t tail0 = head + body_bytes / 4;
        if (threadIdx.x < head) dst[threadIdx.x] = 0.0f;
        if (threadIdx.x < count - tail0) dst[tail0 + threadIdx.x] = 0.0f;
    }
    if (threadIdx.x == 0) {
        const uint64_t pol = l2_policy_evict_first();
        const size_t nchunk = (body_bytes + kFillChunk - 1) / kFillChunk;
        for (size_t c = blockIdx.x; c < nchunk; c += gridDim.x) {
            const size_t off = c * kFillChunk;
            const uint32_t bytes = (uint32_t)((body_bytes - off < kFillChunk) ? body_bytes - off : kFillChunk);
            bulk_store_hint(body + off, zeros, bytes, pol);
            bulk_commit();
            // Bound the bytes this SM has in flight: an unbounded queue of bulk stores delays every other
            // memory request of the kernels running beside the fill.
            if (max_pending <= 2) bulk_wait_pending<2>();
            else if (max_pending <= 4) bulk_wait_pending<4>();
            else if (max_pending <= 8) bulk_wait_pending<8>();
            else if (max_pending <= 16) bulk_wait_pending<16>();
        }
        bulk_wait_all();
    }
}

// ------------------------------------------------------------------------------------------------
// PATCH strategy, reduction: one WARP per tile, persistent warps, tiles taken 32 at a time (lane l reads the bin
// range of tile 32g+l, empty tiles cost nothing more). Per non-empty tile, warp-synchronous:
//   accumulate : as in tile_build_kernel, into the warp's private image (shared-memory atomics)
//   finalise   : touched cells only
//   write      : for each touched cell, the whole 32-byte sectors of the output that overlap it (clipped to the
//                tile), taken from the image -- full sectors, so L2 never fetches a partially written one;
//                everything else in the tile is already zero in memory (fill_tma_kernel)
//   clean      : re-zero the touched cells
// ------------------------------------------------------------------------------------------------
constexpr int kPatchWarps = 8;
__host__ __device__ inline size_t patch_warp_bytes(int tile_cells, int planes)
{
    const size_t img = ((size_t)tile_cells * planes * 4 + 15) & ~(size_t)15;
    return img + (size_t)tile_cells * 8 + ((((size_t)tile_cells + 1) * 4 + 15) & ~(size_t)15);
}

__global__ void __launch_bounds__(kPatchWarps * 32)
warp_patch_kernel(const uint4 *__restrict__ binned, const unsigned *__restrict__ tile_off, FrameOffsets offs,
                  ViewParams p, const float *__restrict__ lut, float *__restrict__ top, int batch)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr unsigned kFull = 0xFFFFFFFFu;
    const int T = p.tile_cells, planes = p.planes, zn = p.zn;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char *mine = smem_raw + (size_t)warp * patch_warp_bytes(T, planes);
    unsigned *img = reinterpret_cast<unsigned *>(mine);
    unsigned long long *key = reinterpret_cast<unsigned long long *>(mine + (((size_t)T * planes * 4 + 15) & ~(size_t)15));
    unsigned *dirty = reinterpret_cast<unsigned *>(key + T);  // dirty[0..T) cells, dirty[T] their number
    for (int w = lane; w < T * planes; w += 32) img[w] = 0u;
    for (int w = lane; w < T; w += 32) key[w] = 0ull;
    if (lane == 0) dirty[T] = 0u;
    __syncwarp();

    const int tpf = p.tiles_per_frame;
    const int groups_per_frame = (tpf + 31) >> 5;
    const int total_groups = batch * groups_per_frame;
    const long long cells_total = (long long)p.xn * p.yn;
    const int nwarps = gridDim.x * kPatchWarps;
    for (int g = blockIdx.x * kPatchWarps + warp; g < total_groups; g += nwarps) {
        const int f = g / groups_per_frame;
        const int t0 = (g - f * groups_per_frame) << 5;
        const unsigned *off = tile_off + (size_t)f * (tpf + 1);
        unsigned beg = 0, end = 0;
        if (t0 + lane < tpf) { beg = __ldg(off + t0 + lane); end = __ldg(off + t0 + lane + 1); }
        unsigned todo = __ballot_sync(kFull, end > beg);
        const uint4 *frame_rec = binned + offs.off[f];
        float *frame_top = top + (size_t)f * p.top_frame_floats;
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const unsigned b = __shfl_sync(kFull, beg, src);
            const int n = (int)(__shfl_sync(kFull, end, src) - b);
            const uint4 *rec = frame_rec + b;
            const long long cell0 = (long long)(t0 + src) * T;
            const int nfl = (int)min((long long)T, cells_total - cell0) * planes;
            float *dst = frame_top + (size_t)cell0 * planes;

            // accumulate
            uint4 q0 = make_uint4(0, 0, 0, 0);
            for (int r = lane; r < n; r += 32) {
                const uint4 q = __ldg(rec + r);
                if (r == lane) q0 = q;
                const unsigned local = q.x & 0xFFFFu, Z = q.x >> 16;
                unsigned *cellw = img + local * planes;
                if (atomicAdd(cellw + zn + 1, 1u) == 0u) dirty[atomicAdd(dirty + T, 1u)] = local;
                if ((int)q.y > 0) {  // h > 0: the reference's maps start at 0 and update on strict '>'
                    atomicMax(cellw + Z, q.y);
                    atomicMax(key + local, ((unsigned long long)q.y << 32) | (unsigned long long)(0xFFFFFFFFu - q.z));
                }
            }
            __syncwarp();
            const int nd = (int)dirty[T];
            // intensity of each cell's winner
            for (int r = lane; r < n; r += 32) {
                const uint4 q = (r == lane) ? q0 : __ldg(rec + r);
                const unsigned local = q.x & 0xFFFFu;
                if ((int)q.y > 0 &&
                    key[local] == (((unsigned long long)q.y << 32) | (unsigned long long)(0xFFFFFFFFu - q.z))) {
                    img[local * planes + zn] = q.w;
                }
            }
            // heights and density of the touched cells (lanes over planes; the intensity slot is not touched here)
            for (int d = 0; d < nd; d++) {
                unsigned *cellw = img + dirty[d] * planes;
                for (int k = lane; k < planes; k += 32) {
                    if (k < zn) {
                        const unsigned raw = cellw[k];
                        if (raw) cellw[k] = __float_as_uint(slice_value(__uint_as_float(raw), p.z_div, k));
                    } else if (k == zn + 1) {
                        const unsigned c = cellw[k];
                        cellw[k] = __float_as_uint(__ldg(lut + (c < 31u ? c : 31u)));
                    }
                }
            }
            __syncwarp();
            // write the sectors that hold touched cells
            const long long abs0 = (long long)(reinterpret_cast<uintptr_t>(dst) >> 2);  // tile start, in floats
            for (int d = 0; d < nd; d++) {
                const long long a = abs0 + (long long)dirty[d] * planes;
                int lo = (int)((a & ~7LL) - abs0), hi = (int)(((a + planes - 1) | 7LL) + 1 - abs0);
                lo = lo < 0 ? 0 : lo;
                hi = hi > nfl ? nfl : hi;
                for (int w = lo + lane; w < hi; w += 32) dst[w] = __uint_as_float(img[w]);
            }
            __syncwarp();
            // clean
            for (int d = 0; d < nd; d++) {
                const unsigned local = dirty[d];
                for (int k = lane; k < planes; k += 32) img[local * planes + k] = 0u;
                if (lane == 0) key[local] = 0ull;
            }
            __syncwarp();
            if (lane == 0) dirty[T] = 0u;
            __syncwarp();
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Front view: write the whole [R][C][F] image from the per-pixel winners, and clear the winners.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kPointThreads)
front_finalize_kernel(const float4 *__restrict__ pts, FrameOffsets offs, ViewParams p,
                      unsigned *__restrict__ fv_winner, float *__restrict__ front)
{
    const int f = blockIdx.y;
    unsigned *winners = fv_winner + (size_t)f * p.front_pixels;
    float *out = front + (size_t)f * p.front_frame_floats;
    for (int pix = blockIdx.x * blockDim.x + threadIdx.x; pix < p.front_pixels; pix += gridDim.x * blockDim.x) {
        const unsigned w = winners[pix];
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

// Pass 1: reduce straight into the (zeroed) output. Height slots hold kRawTag | h bits until finalised,
// the density slot holds the raw count, cellkey holds the (h, ~index) winner key.
__global__ void __launch_bounds__(kPointThreads)
atomic_scatter_kernel(const float4 *__restrict__ pts, FrameOffsets offs, ViewParams p,
                      float *__restrict__ top, unsigned long long *__restrict__ cellkey,
                      unsigned *__restrict__ fv_winner)
{
    const int f = blockIdx.y;
    const int base = offs.off[f];
    const int n = offs.off[f + 1] - base;
    unsigned *out = reinterpret_cast<unsigned *>(top + (size_t)f * p.top_frame_floats);
    unsigned long long *keys = cellkey + (size_t)f * ((size_t)p.xn * p.yn);
    unsigned *winners = fv_winner + (size_t)f * p.front_pixels;
    // two points per thread and round: both loads are issued before the first atomic
    const int stride = gridDim.x * blockDim.x;
    for (int i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += 2 * stride) {
        float4 pt[2];
        pt[0] = load_point(pts + base + i0);
        if (i0 + stride < n) pt[1] = load_point(pts + base + i0 + stride);
#pragma unroll
        for (int k = 0; k < 2; k++) {
            const int i = i0 + k * stride;
            if (i >= n) continue;
            const CellHit c = classify(pt[k], p);
            if (!c.keep) continue;
            if (p.want_top) {
                unsigned *cellw = out + (size_t)c.cell * p.planes;
                atomicAdd(cellw + p.zn + 1, 1u);
                if (c.h > 0.0f) {
                    atomicMax(cellw + c.slice, __float_as_uint(c.h) | kRawTag);
                    atomicMax(keys + c.cell, cell_key(c.h, (uint32_t)i));
                }
            }
            if (p.want_front) {
                const int pix = front_pixel(pt[k], p);
                if (pix >= 0) atomicMax(winners + pix, (unsigned)i + 1u);
            }
        }
    }
}

// Pass 2: every kept point revisits its cell. The slot winner (tagged value == its own) converts the height;
// the key winner writes the intensity; anyone who still sees a raw count converts it.
// Raw counts (< 2^30) and density floats (>= 0.2f = 0x3E4CCCCD) cannot be confused, and all writers of one
// density slot write the same value, so the race between them is benign.
__global__ void __launch_bounds__(kPointThreads)
atomic_finalize_kernel(const float4 *__restrict__ pts, FrameOffsets offs, ViewParams p, const float *__restrict__ lut,
                       float *__restrict__ top, unsigned long long *__restrict__ cellkey)
{
    const int f = blockIdx.y;
    const int base = offs.off[f];
    const int n = offs.off[f + 1] - base;
    unsigned *out = reinterpret_cast<unsigned *>(top + (size_t)f * p.top_frame_floats);
    unsigned long long *keys = cellkey + (size_t)f * ((size_t)p.xn * p.yn);
    const int stride = gridDim.x * blockDim.x;
    for (int i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += 2 * stride) {
        float4 pt[2];
        pt[0] = load_point(pts + base + i0);
        if (i0 + stride < n) pt[1] = load_point(pts + base + i0 + stride);
        // both points' cell words are requested before either is examined
        CellHit c[2];
        unsigned cnt[2] = {0u, 0u}, slot[2] = {0u, 0u};
        unsigned long long key[2] = {0ull, 0ull};
#pragma unroll
        for (int k = 0; k < 2; k++) {
            c[k].keep = false;
            if (i0 + k * stride >= n) continue;
            c[k] = classify(pt[k], p);
            if (!c[k].keep) continue;
            unsigned *cellw = out + (size_t)c[k].cell * p.planes;
            cnt[k] = *reinterpret_cast<volatile unsigned *>(cellw + p.zn + 1);
            if (c[k].h > 0.0f) {
                slot[k] = *reinterpret_cast<volatile unsigned *>(cellw + c[k].slice);
                key[k] = *reinterpret_cast<volatile unsigned long long *>(keys + c[k].cell);
            }
        }
#pragma unroll
        for (int k = 0; k < 2; k++) {
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

// ------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------
static inline int point_blocks(int max_points, int items = 1)
{
    int b = (max_points + kPointThreads * items - 1) / (kPointThreads * items);
    return b < 1 ? 1 : (b > 4096 ? 4096 : b);
}

size_t tile_smem_bytes(const ViewParams &p)
{
    const size_t img = ((size_t)p.tile_cells * p.planes * 4 + 15) & ~(size_t)15;
    return img + (size_t)p.tile_cells * 8 + (size_t)p.tile_cells * 4 + ((size_t)p.tile_cells / 32 + 2) * 4;
}

cudaError_t tile_build_configure(const ViewParams &p, int *blocks_per_sm)
{
    const size_t smem = tile_smem_bytes(p);
    auto kernel = p.py ? tile_build_kernel<true> : tile_build_kernel<false>;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, kernel, kTileThreads, smem);
}

int launch_classify(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, int batch, int max_points,
                    uint2 *keys, unsigned *tile_count, unsigned *fv_winner, cudaStream_t s)
{
    dim3 grid(point_blocks(max_points, kPointItems), batch);
    auto kernel = p.py ? classify_kernel<true> : classify_kernel<false>;
    kernel<<<grid, kPointThreads, 0, s>>>(pts, offs, p, keys, tile_count, fv_winner);
    return 1;
}

int launch_tile_scan(unsigned *tile_count, unsigned *tile_off, int tiles_per_frame, int batch, cudaStream_t s)
{
    if (tiles_per_frame > 16 * kScanThreads)
        tile_scan_kernel<8><<<batch, kScanThreads, 0, s>>>(tile_count, tile_off, tiles_per_frame);
    else
        tile_scan_kernel<2><<<batch, kScanThreads, 0, s>>>(tile_count, tile_off, tiles_per_frame);
    return 1;
}

int launch_bin(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, int batch, int max_points,
               const uint2 *keys, const unsigned *tile_off, uint4 *binned, cudaStream_t s)
{
    dim3 grid(point_blocks(max_points, kPointItems), batch);
    auto kernel = p.py ? bin_kernel<true> : bin_kernel<false>;
    kernel<<<grid, kPointThreads, 0, s>>>(pts, offs, p, keys, tile_off, binned);
    return 1;
}

int launch_tile_build(const uint4 *binned, const unsigned *tile_off, const FrameOffsets &offs, const ViewParams &p,
                      const float *lut, float *top, int batch, int grid_blocks, cudaStream_t s, const OccupancySink *sink)
{
    const int total = batch * p.tiles_per_frame;
    if (p.pipe_groups > 0 && (long long)total + 2LL * p.pipe_groups * p.pipe_grid < 2147483647LL) {  // one CTA per SM
        const int grid = p.pipe_grid < total ? p.pipe_grid : total;
        const bool occ = sink && sink->words;
        auto kernel = occ ? (p.py ? tile_pipe_kernel<true, true> : tile_pipe_kernel<true, false>)
                          : (p.py ? tile_pipe_kernel<false, true> : tile_pipe_kernel<false, false>);
        const size_t smem = tile_pipe_smem_bytes(p);
        cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        const int threads = p.pipe_groups * p.pipe_group_threads + 32;
        kernel<<<grid, threads, smem, s>>>(binned, tile_off, offs, p, lut, top, total, occ ? *sink : OccupancySink());
        return 1;
    }
    const int grid = grid_blocks < total ? grid_blocks : total;
    if (sink && sink->words) {
        auto kernel = p.py ? tile_build_occupancy_kernel<true> : tile_build_occupancy_kernel<false>;
        cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tile_smem_bytes(p));
        kernel<<<grid, kTileThreads, tile_smem_bytes(p), s>>>(binned, tile_off, offs, p, lut, top, total, *sink);
    } else {
        auto kernel = p.py ? tile_build_kernel<true> : tile_build_kernel<false>;
        if (!p.py && p.store_mode != kStoreTmaHint) {  // experiment knob (MV3D_TILE_STORE), C-routine semantics only
            kernel = p.store_mode == kStoreTma ? tile_build_kernel<false, kStoreTma>
                   : p.store_mode == kStoreTmaSplit ? tile_build_kernel<false, kStoreTmaSplit>
                   : p.store_mode == kStoreTmaPaced ? tile_build_kernel<false, kStoreTmaPaced>
                   : p.store_mode == kStoreLsu ? tile_build_kernel<false, kStoreLsu>
                                               : tile_build_kernel<false, kStoreLsuStream>;
            cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tile_smem_bytes(p));
        }
        kernel<<<grid, kTileThreads, tile_smem_bytes(p), s>>>(binned, tile_off, offs, p, lut, top, total);
    }
    return 1;
}

cudaError_t tile_patch_configure(const ViewParams &p, int *blocks_per_sm)
{
    const size_t smem = tile_smem_bytes(p);
    auto kernel = p.py ? tile_patch_kernel<true> : tile_patch_kernel<false>;
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, kernel, kTileThreads, smem);
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
                        cudaStream_t s)
{
    const int total = batch * p.tiles_per_frame;
    const int grid = grid_blocks < total ? grid_blocks : total;
    auto kernel = p.py ? tile_compact_kernel<true> : tile_compact_kernel<false>;
    cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tile_smem_bytes(p));
    kernel<<<grid, kTileThreads, tile_smem_bytes(p), s>>>(binned, tile_off, offs, p, lut, total, compact, compact_cap);
    return 1;
}

size_t patch_smem_bytes(const ViewParams &p) { return kPatchWarps * patch_warp_bytes(p.tile_cells, p.planes); }

cudaError_t warp_patch_configure(const ViewParams &p, int *blocks_per_sm)
{
    const size_t smem = patch_smem_bytes(p);
    cudaError_t e = cudaFuncSetAttribute(warp_patch_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, warp_patch_kernel, kPatchWarps * 32, smem);
}

size_t fill_tma_smem_bytes() { return kFillChunk; }

int launch_fill_tma(float *dst, size_t count, int blocks, int max_pending, cudaStream_t s)
{
    if (count == 0) return 0;
    const size_t want = (count * 4 + kFillChunk - 1) / kFillChunk;
    if ((size_t)blocks > want) blocks = (int)want;
    fill_tma_kernel<<<blocks < 1 ? 1 : blocks, kFillThreads, 0, s>>>(dst, count, max_pending);
    return 1;
}

int launch_warp_patch(const uint4 *binned, const unsigned *tile_off, const FrameOffsets &offs, const ViewParams &p,
                      const float *lut, float *top, int batch, int grid_blocks, cudaStream_t s)
{
    const int groups = batch * ((p.tiles_per_frame + 31) / 32);
    const int want = (groups + kPatchWarps - 1) / kPatchWarps;
    const int grid = grid_blocks < want ? grid_blocks : want;
    warp_patch_kernel<<<grid < 1 ? 1 : grid, kPatchWarps * 32, patch_smem_bytes(p), s>>>(binned, tile_off, offs, p, lut, top, batch);
    return 1;
}

int launch_front_finalize(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, int batch,
                          unsigned *fv_winner, float *front, cudaStream_t s)
{
    dim3 grid(point_blocks(p.front_pixels), batch);
    front_finalize_kernel<<<grid, kPointThreads, 0, s>>>(pts, offs, p, fv_winner, front);
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
    atomic_scatter_kernel<<<grid, kPointThreads, 0, s>>>(pts, offs, p, top, cellkey, fv_winner);
    return 1;
}

int launch_atomic_finalize(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, const float *lut,
                           int batch, int max_points, float *top, unsigned long long *cellkey, cudaStream_t s)
{
    dim3 grid(point_blocks(max_points), batch);
    atomic_finalize_kernel<<<grid, kPointThreads, 0, s>>>(pts, offs, p, lut, top, cellkey);
    return 1;
}

} // namespace mv3d
