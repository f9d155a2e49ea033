// anchor_kernels.cu -- the empty-anchor filter on a device-resident bird's-eye view, for sm_100a.
//
// What it replaces: src/net/utility/remove_empty_box_kernel.cu:3-42 launched by
// src/net/utility/remove_empty_box.py:25-47 (one block per anchor row, an unsynchronised `+=` per channel, the view
// uploaded on every call). Semantics: include/mv3d_anchors.h.
//
// Two ways to the same decision, chosen per frame on the device:
//   anchor_occupancy_kernel  one streaming read of the view [cells][planes] (HBM-bound, planes*4 bytes per cell):
//                            one bit per cell "holds a non-zero value", and a per-frame flag "holds a negative
//                            value, a -0 or a NaN". Without the flag every value is >= +0, float32 sums of such
//                            values are monotone, and "total > 0" is exactly "some cell of the box has its bit set".
//   anchor_rows_kernel       the bit map re-laid row by row (word-aligned rows) plus its OR over groups of 8 rows
//   anchor_test_kernel       thread per anchor: ORs the column range of the box over its rows, 8 rows at a time where
//                            a whole group lies inside (<= 50 rows -> <= 20 probes of 1-3 words).
//   anchor_sum_kernel        frames with the flag, limit != 0, or totals requested: warp per anchor, lane per
//                            channel, the float32 sums in the reference's order (row left to right, rows ascending,
//                            numpy's pairwise sum over the channels), read through L2 (a 52 MB view fits it).
//   anchor_compact_kernel    keep flags -> ascending index list (np.where order) and count, per frame.
#include "anchor_kernels.h"

#include <algorithm>

namespace mv3d {

namespace {

constexpr unsigned kFull = 0xffffffffu;

__device__ __forceinline__ void note_value(unsigned bits, unsigned cell, unsigned &mask, unsigned &orbits,
                                           unsigned &maxabs)
{
    mask |= (bits != 0u ? 1u : 0u) << cell;
    orbits |= bits;
    maxabs = max(maxabs, bits & 0x7fffffffu);
}

// One warp per occupancy word = 32 consecutive cells = 32*planes consecutive floats (128*planes bytes: 16-byte
// aligned whenever the frame is). `magic` = ceil(2^32 / planes) turns a float offset inside the word into its cell.
template <bool kVec>
__global__ void __launch_bounds__(256) anchor_occupancy_kernel(const float *__restrict__ views, size_t frame_floats,
                                                               unsigned cells, unsigned planes, unsigned magic,
                                                               unsigned words, unsigned *__restrict__ occupancy,
                                                               unsigned *__restrict__ flags)
{
    const unsigned lane = threadIdx.x & 31u;
    const unsigned warps = gridDim.x * (blockDim.x >> 5);
    const unsigned frame = blockIdx.y;
    const float *view = views + (size_t)frame * frame_floats;
    for (unsigned w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); w < words; w += warps) {
        const unsigned ncell = min(32u, cells - 32u * w);
        const unsigned count = ncell * planes;
        const float *base = view + (size_t)w * 32u * planes;
        unsigned mask = 0, orbits = 0, maxabs = 0;
        if (kVec) {
            const uint4 *base4 = reinterpret_cast<const uint4 *>(base);
            const unsigned n4 = count >> 2;
            for (unsigned q0 = 0; q0 < n4; q0 += 128u) {
                uint4 v[4];
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const unsigned q = q0 + 32u * k + lane;
                    v[k] = q < n4 ? __ldg(base4 + q) : make_uint4(0, 0, 0, 0);
                }
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const unsigned e = 4u * (q0 + 32u * k + lane);  // out-of-range slots hold zeros: any cell will do
                    const unsigned c0 = planes == 1 ? e : __umulhi(e, magic);
                    const unsigned c1 = planes == 1 ? e + 1 : __umulhi(e + 1, magic);
                    const unsigned c2 = planes == 1 ? e + 2 : __umulhi(e + 2, magic);
                    const unsigned c3 = planes == 1 ? e + 3 : __umulhi(e + 3, magic);
                    note_value(v[k].x, c0 & 31u, mask, orbits, maxabs);
                    note_value(v[k].y, c1 & 31u, mask, orbits, maxabs);
                    note_value(v[k].z, c2 & 31u, mask, orbits, maxabs);
                    note_value(v[k].w, c3 & 31u, mask, orbits, maxabs);
                }
            }
            for (unsigned e = (n4 << 2) + lane; e < count; e += 32u) {
                const unsigned c = planes == 1 ? e : __umulhi(e, magic);
                note_value(__float_as_uint(__ldg(base + e)), c, mask, orbits, maxabs);
            }
        } else {
            for (unsigned e = lane; e < count; e += 32u) {
                const unsigned c = planes == 1 ? e : __umulhi(e, magic);
                note_value(__float_as_uint(__ldg(base + e)), c, mask, orbits, maxabs);
            }
        }
        mask = __reduce_or_sync(kFull, mask);
        orbits = __reduce_or_sync(kFull, orbits);
        maxabs = __reduce_max_sync(kFull, maxabs);
        if (lane == 0) {
            occupancy[(size_t)frame * words + w] = mask;
            if ((orbits >> 31) != 0u || maxabs > 0x7f800000u) atomicOr(flags + frame, 1u);
        }
    }
}

struct Box {
    int x1, y1, lines, len;
};

// remove_empty_box_kernel.cu:16-33: (y1, x1, y2, x2), each clamped to [0, dim-1]; rows [x1, x2), columns [y1, y2).
__device__ __forceinline__ Box clamp_box(int4 b, int xn, int yn, int max_lines)
{
    const int y1 = min(max(b.x, 0), yn - 1), x1 = min(max(b.y, 0), xn - 1);
    const int y2 = min(max(b.z, 0), yn - 1), x2 = min(max(b.w, 0), xn - 1);
    Box r;
    r.x1 = x1;
    r.y1 = y1;
    r.lines = min(x2 - x1, max_lines);
    r.len = y2 - y1;
    return r;
}

// Row-aligned copy of the occupancy map, and its OR over groups of 8 rows. The map is laid out over the flattened
// cell index (that is how the view is stored and how tiles cut it), so a row of the view starts at an arbitrary bit;
// here row x starts at word x * wpr, which makes the column masks of a box the same for all of its rows and lets the
// test take 8 rows at a time.  thread = (group of 8 rows, word column).
__global__ void __launch_bounds__(256) anchor_rows_kernel(AnchorGeometry g, const unsigned *__restrict__ occupancy,
                                                          const unsigned *__restrict__ flags,
                                                          unsigned *__restrict__ rows, unsigned *__restrict__ groups)
{
    const unsigned frame = blockIdx.y;
    if (flags[frame] != 0u) return;
    const unsigned wpr = g.wpr, ngroups = (unsigned)(g.xn + 7) >> 3;
    const unsigned item = blockIdx.x * blockDim.x + threadIdx.x;
    if (item >= ngroups * wpr) return;
    const unsigned grp = item / wpr, w = item - grp * wpr;
    const unsigned *occ = occupancy + (size_t)frame * g.words;
    const unsigned nbits = min(32u, (unsigned)g.yn - 32u * w);
    const unsigned keep = nbits == 32u ? kFull : ((1u << nbits) - 1u);
    unsigned any = 0;
    for (unsigned x = grp * 8u; x < min(grp * 8u + 8u, (unsigned)g.xn); x++) {
        const unsigned b = x * (unsigned)g.yn + 32u * w;
        const unsigned i = b >> 5;
        const unsigned lo = __ldg(occ + i), hi = (i + 1 < g.words) ? __ldg(occ + i + 1) : 0u;
        const unsigned word = __funnelshift_r(lo, hi, b & 31u) & keep;
        rows[((size_t)frame * g.xn + x) * wpr + w] = word;
        any |= word;
    }
    groups[((size_t)frame * ngroups + grp) * wpr + w] = any;
}

__global__ void __launch_bounds__(kAnchorChunk) anchor_test_kernel(const int4 *__restrict__ boxes, AnchorGeometry g,
                                                                   const unsigned *__restrict__ rows,
                                                                   const unsigned *__restrict__ groups,
                                                                   const unsigned *__restrict__ flags,
                                                                   unsigned char *__restrict__ keep,
                                                                   unsigned *__restrict__ chunk_count)
{
    const unsigned frame = blockIdx.y;
    if (flags[frame] != 0u) return;  // this frame needs the exact sums (anchor_sum_kernel takes it)
    const int a = blockIdx.x * kAnchorChunk + threadIdx.x;
    bool hit = false;
    if (a < g.n) {
        const Box b = clamp_box(boxes[a], g.xn, g.yn, g.max_lines);
        if (b.lines > 0 && b.len > 0) {
            const unsigned wpr = g.wpr;
            const unsigned y_last = (unsigned)(b.y1 + b.len - 1);
            const unsigned w0 = (unsigned)b.y1 >> 5, w1 = y_last >> 5;
            const unsigned m0 = kFull << ((unsigned)b.y1 & 31u), m1 = kFull >> (31u - (y_last & 31u));
            auto probe = [&](const unsigned *row) {  // any bit of columns [y1, y1 + len) in this row (or 8-row group)?
                unsigned any;
                if (w0 == w1) {
                    any = __ldg(row + w0) & m0 & m1;
                } else {
                    any = (__ldg(row + w0) & m0) | (__ldg(row + w1) & m1);
                    for (unsigned w = w0 + 1; w < w1; w++) any |= __ldg(row + w);
                }
                return any != 0u;
            };
            const unsigned *frame_rows = rows + (size_t)frame * g.xn * wpr;
            const unsigned *frame_groups = groups + (size_t)frame * ((unsigned)(g.xn + 7) >> 3) * wpr;
            int x = b.x1;
            const int xend = b.x1 + b.lines;
            while (!hit && x < xend && (x & 7)) hit = probe(frame_rows + (size_t)x++ * wpr);        // up to a group boundary
            while (!hit && x + 8 <= xend) { hit = probe(frame_groups + (size_t)(x >> 3) * wpr); x += 8; }  // whole groups
            while (!hit && x < xend) hit = probe(frame_rows + (size_t)x++ * wpr);                    // the rest
        }
        keep[(size_t)frame * g.n + a] = hit ? 1 : 0;
    }
    const int kept = __syncthreads_count(hit ? 1 : 0);
    if (threadIdx.x == 0) chunk_count[(size_t)frame * g.chunks + blockIdx.x] = (unsigned)kept;
}

// numpy's float32 add-reduce of a contiguous row (the `np.sum(index, axis=1)` of remove_empty_box.py:46):
// 0 + pairwise(v): fewer than 8 values -> a running sum from -0; up to 128 -> eight running sums combined as
// ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)) then the tail; longer rows split in two at a multiple of 8.
__device__ float pairwise_sum(const float *v, int n)
{
    if (n < 8) {
        float r = -0.0f;
        for (int i = 0; i < n; i++) r = __fadd_rn(r, v[i]);
        return r;
    }
    if (n <= 128) {
        float r[8];
#pragma unroll
        for (int j = 0; j < 8; j++) r[j] = v[j];
        int i = 8;
        for (; i < n - (n % 8); i += 8) {
#pragma unroll
            for (int j = 0; j < 8; j++) r[j] = __fadd_rn(r[j], v[i + j]);
        }
        float res = __fadd_rn(__fadd_rn(__fadd_rn(r[0], r[1]), __fadd_rn(r[2], r[3])),
                              __fadd_rn(__fadd_rn(r[4], r[5]), __fadd_rn(r[6], r[7])));
        for (; i < n; i++) res = __fadd_rn(res, v[i]);
        return res;
    }
    int n2 = n / 2;
    n2 -= n2 % 8;
    return __fadd_rn(pairwise_sum(v, n2), pairwise_sum(v + n2, n - n2));
}

__global__ void __launch_bounds__(kAnchorSumWarps * 32) anchor_sum_kernel(const float *__restrict__ views,
                                                                          const int4 *__restrict__ boxes,
                                                                          AnchorGeometry g, float limit,
                                                                          const unsigned *__restrict__ flags,
                                                                          int always, unsigned char *__restrict__ keep,
                                                                          unsigned *__restrict__ chunk_count,
                                                                          float *__restrict__ totals)
{
    extern __shared__ float channel_sums[];  // [kAnchorSumWarps][planes]
    __shared__ int kept[kAnchorSumWarps];
    const unsigned frame = blockIdx.y;
    if (!always && flags[frame] == 0u) return;  // decided by the occupancy map
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    float *sums = channel_sums + (size_t)warp * g.planes;
    const int groups = (g.n + kAnchorSumWarps - 1) / kAnchorSumWarps;
    for (int group = blockIdx.x; group < groups; group += gridDim.x) {
    const int a = group * kAnchorSumWarps + warp;
    int hit = 0;
    if (a < g.n) {
        const Box b = clamp_box(boxes[a], g.xn, g.yn, g.max_lines);
        const float *view = views + (size_t)frame * g.cells * g.planes;
        for (int ch = lane; ch < g.planes; ch += 32) {
            float acc = 0.0f;  // inds[] starts at zero (remove_empty_box.py:35)
            for (int line = 0; line < b.lines; line++) {
                const float *row = view + ((size_t)(b.x1 + line) * g.yn + b.y1) * g.planes + ch;
                float sum = 0.0f;  // `integration`, remove_empty_box_kernel.cu:3-10
#pragma unroll 8
                for (int i = 0; i < b.len; i++) sum = __fadd_rn(sum, __ldg(row + (size_t)i * g.planes));
                acc = __fadd_rn(acc, sum);  // :41, rows in ascending order
            }
            sums[ch] = acc;
        }
        __syncwarp();
        if (lane == 0) {
            const float total = __fadd_rn(0.0f, pairwise_sum(sums, g.planes));
            hit = total > limit ? 1 : 0;
            keep[(size_t)frame * g.n + a] = (unsigned char)hit;
            if (totals) totals[(size_t)frame * g.n + a] = total;
        }
    }
    if (lane == 0) kept[warp] = hit;
    __syncthreads();
    if (threadIdx.x == 0) {
        int sum = 0;
#pragma unroll
        for (int w = 0; w < kAnchorSumWarps; w++) sum += kept[w];
        // the anchors of one group share a chunk: kAnchorChunk is a multiple of kAnchorSumWarps
        if (sum) atomicAdd(chunk_count + (size_t)frame * g.chunks + (group * kAnchorSumWarps) / kAnchorChunk, sum);
    }
    __syncthreads();  // kept[] is reused by the next group
    }
}

__global__ void __launch_bounds__(kAnchorChunk) anchor_compact_kernel(AnchorGeometry g,
                                                                      const unsigned char *__restrict__ keep,
                                                                      const unsigned *__restrict__ chunk_count,
                                                                      int *__restrict__ keep_idx, int *__restrict__ count)
{
    __shared__ unsigned warp_part[32];
    __shared__ unsigned base_shared, total_shared;
    const unsigned frame = blockIdx.y, blk = blockIdx.x;
    const unsigned tid = threadIdx.x, warp = tid >> 5, lane = tid & 31u;
    const unsigned *cc = chunk_count + (size_t)frame * g.chunks;
    // kept anchors in the chunks before this one
    unsigned part = 0;
    for (unsigned c = tid; c < blk; c += kAnchorChunk) part += cc[c];
    part = __reduce_add_sync(kFull, part);
    if (lane == 0) warp_part[warp] = part;
    __syncthreads();
    if (warp == 0) {
        const unsigned v = __reduce_add_sync(kFull, warp_part[lane]);
        if (lane == 0) base_shared = v;
    }
    __syncthreads();
    const int a = (int)(blk * kAnchorChunk + tid);
    const bool flag = a < g.n && keep[(size_t)frame * g.n + a] != 0;
    const unsigned ballot = __ballot_sync(kFull, flag);
    const unsigned rank = __popc(ballot & ((1u << lane) - 1u));
    if (lane == 0) warp_part[warp] = __popc(ballot);
    __syncthreads();
    if (warp == 0) {
        const unsigned v = warp_part[lane];
        unsigned incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned up = __shfl_up_sync(kFull, incl, d);
            if ((int)lane >= d) incl += up;
        }
        warp_part[lane] = incl - v;
        if (lane == 31) total_shared = incl;
    }
    __syncthreads();
    if (flag) keep_idx[(size_t)frame * g.n + base_shared + warp_part[warp] + rank] = a;
    if (blk == g.chunks - 1 && tid == 0) count[frame] = (int)(base_shared + total_shared);
}

} // namespace

int launch_anchor_occupancy(const float *views, const AnchorGeometry &g, int frames, unsigned *occupancy,
                            unsigned *flags, int sm_count, cudaStream_t s)
{
    const size_t frame_floats = (size_t)g.cells * g.planes;
    const unsigned magic = g.planes == 1 ? 0u : 0xffffffffu / (unsigned)g.planes + 1u;
    const bool vec = (reinterpret_cast<uintptr_t>(views) & 15u) == 0 && frame_floats % 4 == 0;
    // 8 warps per block, a word per warp at a time; at most ~4 waves of resident blocks over all frames
    unsigned blocks = (g.words + 7u) / 8u;
    const unsigned cap = std::max(1u, (unsigned)sm_count * 32u / (unsigned)frames);
    if (blocks > cap) blocks = cap;
    dim3 grid(blocks, frames);
    if (vec)
        anchor_occupancy_kernel<true><<<grid, 256, 0, s>>>(views, frame_floats, g.cells, g.planes, magic, g.words,
                                                          occupancy, flags);
    else
        anchor_occupancy_kernel<false><<<grid, 256, 0, s>>>(views, frame_floats, g.cells, g.planes, magic, g.words,
                                                           occupancy, flags);
    return 1;
}

int launch_anchor_test(const int4 *boxes, const AnchorGeometry &g, int frames, const unsigned *occupancy,
                       const unsigned *flags, unsigned *rows, unsigned *groups, unsigned char *keep,
                       unsigned *chunk_count, cudaStream_t s)
{
    const unsigned items = ((unsigned)(g.xn + 7) >> 3) * g.wpr;
    anchor_rows_kernel<<<dim3((items + 255u) / 256u, frames), 256, 0, s>>>(g, occupancy, flags, rows, groups);
    anchor_test_kernel<<<dim3(g.chunks, frames), kAnchorChunk, 0, s>>>(boxes, g, rows, groups, flags, keep, chunk_count);
    return 2;
}

int launch_anchor_sum(const float *views, const int4 *boxes, const AnchorGeometry &g, int frames, float limit,
                      const unsigned *flags, bool always, unsigned char *keep, unsigned *chunk_count, float *totals,
                      int sm_count, cudaStream_t s)
{
    // Blocks walk the anchors 8 at a time. A grid of about four waves over all frames keeps the launch that finds
    // nothing to do (every frame decided by the occupancy map) down to a few microseconds.
    unsigned blocks = (g.n + kAnchorSumWarps - 1) / kAnchorSumWarps;
    const unsigned cap = std::max(1u, (unsigned)sm_count * 32u / (unsigned)frames);
    if (blocks > cap) blocks = cap;
    const size_t smem = (size_t)kAnchorSumWarps * g.planes * sizeof(float);
    anchor_sum_kernel<<<dim3(blocks, frames), kAnchorSumWarps * 32, smem, s>>>(views, boxes, g, limit, flags,
                                                                               always ? 1 : 0, keep, chunk_count, totals);
    return 1;
}

int launch_anchor_compact(const AnchorGeometry &g, int frames, const unsigned char *keep, const unsigned *chunk_count,
                          int *keep_idx, int *count, cudaStream_t s)
{
    anchor_compact_kernel<<<dim3(g.chunks, frames), kAnchorChunk, 0, s>>>(g, keep, chunk_count, keep_idx, count);
    return 1;
}

} // namespace mv3d
