// sort_strategy.cu -- the SORT strategy of the bird's-eye view: sort the kept points by (frame, cell, slice) and
// reduce each cell's run of consecutive entries. It exists as the measured alternative the choice of strategy is made
// against (BASELINE north star: "shared-memory-privatised tiles ... or a radix sort by cell key followed by a
// segmented reduction"); see profiles/strategy_experiments_r1.md for the numbers.
//
//   sort_keys     one pass over the points (coalesced 16-byte loads, same arithmetic as classify): 32-bit key
//                 (frame * cells + cell) * Zn + slice, value = index of the point in its sweep; rejected points get
//                 the key one past the largest, so they sort to the end. Elects the front-view winners on the way.
//   radix sort    cub::DeviceRadixSort::SortPairs over the bits the keys need -- a LIBRARY kernel (CCCL ships with the
//                 CUDA toolkit). It is stable, so equal keys keep input order.
//   fill_zero     the output is written dense, zeros included.
//   sort_reduce   one thread per sorted entry stages (cell, slice, height, intensity, index) in shared memory with
//                 coalesced loads and a gather of the points; the thread at the head of a cell's run then walks the
//                 run (out of shared memory while it stays inside the CTA's 256 entries, from global memory beyond),
//                 keeps the running slice maximum, the count and the (height, ~index) winner, and writes the finished
//                 values of the cell. No atomics, no second pass: every touched value is written once.
// C-routine semantics only (the Python twin's two-slice membership needs two keys per point).
#include <cub/device/device_radix_sort.cuh>

#include "encode_kernels.h"
#include "view_math.cuh"
#include "selfcheck.cuh"

namespace mv3d {

constexpr int kSortThreads = 256;

__global__ void __launch_bounds__(kSortThreads)
sort_keys_kernel(const float4 *__restrict__ pts, const __grid_constant__ FrameOffsets offs,
                 const __grid_constant__ ViewParams p, unsigned *__restrict__ keys, unsigned *__restrict__ vals,
                 unsigned *__restrict__ fv_winner, unsigned invalid_key)
{
    const int f = blockIdx.y;
    const int base = offs.off[f];
    const int n = offs.off[f + 1] - base;
    unsigned *winners = fv_winner + (size_t)f * p.front_pixels;
    const unsigned cells = (unsigned)(p.xn * p.yn);
    const int stride = gridDim.x * blockDim.x;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const float4 pt = load_point(pts + base + i);
        const CellHit c = classify_as<false>(pt, p);
        unsigned key = invalid_key;
        if (c.keep) {
            key = ((unsigned)f * cells + c.cell) * (unsigned)p.zn + (unsigned)c.slice;
            if (p.want_front) {
                const int pix = front_pixel(pt, p);
                if (pix >= 0) atomicMax(winners + pix, (unsigned)i + 1u);
            }
        }
        keys[base + i] = key;
        vals[base + i] = (unsigned)i;
    }
}

struct SortedEntry {
    unsigned cell;  // frame * cells + cell, or 0xFFFFFFFF
    unsigned slice, hbits, wbits, index;
};

__device__ __forceinline__ SortedEntry fetch_entry(const float4 *__restrict__ pts, const FrameOffsets &offs,
                                                   const ViewParams &p, const unsigned *__restrict__ skeys,
                                                   const unsigned *__restrict__ svals, size_t i, unsigned invalid_key,
                                                   unsigned cells)
{
    SortedEntry e;
    const unsigned key = skeys[i];
    e.cell = 0xFFFFFFFFu;
    e.slice = e.hbits = e.wbits = e.index = 0u;
    if (key < invalid_key) {
        e.cell = key / (unsigned)p.zn;
        e.slice = key - e.cell * (unsigned)p.zn;
        e.index = svals[i];
        const unsigned f = e.cell / cells;
        MV3D_CHECK(f < (unsigned)kMaxBatchPerLaunch && (int)e.index < offs.off[f + 1] - offs.off[f], MV3D_BAD_WINNER);
        const float4 pt = __ldg(pts + offs.off[f] + e.index);
        e.hbits = __float_as_uint(__fsub_rn(pt.z, p.z_min));
        e.wbits = __float_as_uint(pt.w);
    }
    return e;
}

__global__ void __launch_bounds__(kSortThreads)
sort_reduce_kernel(const float4 *__restrict__ pts, const __grid_constant__ FrameOffsets offs,
                   const __grid_constant__ ViewParams p, const float *__restrict__ lut,
                   const unsigned *__restrict__ skeys, const unsigned *__restrict__ svals, size_t n,
                   unsigned invalid_key, float *__restrict__ top)
{
    __shared__ unsigned s_cell[kSortThreads], s_slice[kSortThreads], s_h[kSortThreads], s_w[kSortThreads],
        s_idx[kSortThreads];
    const unsigned cells = (unsigned)(p.xn * p.yn);
    const int tid = threadIdx.x;
    const size_t block0 = (size_t)blockIdx.x * kSortThreads;
    const size_t i = block0 + tid;
    SortedEntry e;
    e.cell = 0xFFFFFFFFu;
    e.slice = e.hbits = e.wbits = e.index = 0u;
    if (i < n) e = fetch_entry(pts, offs, p, skeys, svals, i, invalid_key, cells);
    s_cell[tid] = e.cell; s_slice[tid] = e.slice; s_h[tid] = e.hbits; s_w[tid] = e.wbits; s_idx[tid] = e.index;
    __syncthreads();
    if (e.cell == 0xFFFFFFFFu) return;
    unsigned prev = 0xFFFFFFFFu;
    if (tid > 0) prev = s_cell[tid - 1];
    else if (i > 0) {
        const unsigned pk = skeys[i - 1];
        if (pk < invalid_key) prev = pk / (unsigned)p.zn;
    }
    if (prev == e.cell) return;  // not the head of its cell's run

    MV3D_CHECK(e.slice < (unsigned)p.zn, MV3D_BAD_SLICE);
    float *out = top + (size_t)e.cell * p.planes;  // frames are contiguous: (frame * cells + cell) * planes
    unsigned count = 0, cur_slice = e.slice, cur_max = 0u, best_w = 0u;
    unsigned long long best = 0ull;
    SortedEntry r = e;
    size_t j = i;
    for (;;) {
        count++;
        if (r.slice != cur_slice) {
            if (cur_max) out[cur_slice] = slice_value(__uint_as_float(cur_max), p.z_div, (int)cur_slice);
            cur_slice = r.slice;
            cur_max = 0u;
        }
        if ((int)r.hbits > 0) {  // h > 0: the reference's maps start at 0 and update on strict '>'
            if (r.hbits > cur_max) cur_max = r.hbits;
            const unsigned long long k = ((unsigned long long)r.hbits << 32) | (unsigned long long)(0xFFFFFFFFu - r.index);
            if (k > best) { best = k; best_w = r.wbits; }
        }
        j++;
        if (j >= n) break;
        const size_t local = j - block0;
        if (local < (size_t)kSortThreads) {
            r.cell = s_cell[local]; r.slice = s_slice[local]; r.hbits = s_h[local]; r.wbits = s_w[local]; r.index = s_idx[local];
        } else {
            r = fetch_entry(pts, offs, p, skeys, svals, j, invalid_key, cells);
        }
        if (r.cell != e.cell) break;
    }
    if (cur_max) out[cur_slice] = slice_value(__uint_as_float(cur_max), p.z_div, (int)cur_slice);
    if (best) out[p.zn] = __uint_as_float(best_w);
    out[p.zn + 1] = __ldg(lut + (count < 31u ? count : 31u));
}

// Largest key + 1 (the "rejected" key) for this call, or 0 when the keys do not fit 32 bits.
unsigned sort_invalid_key(const ViewParams &p, int batch)
{
    const unsigned long long zn = p.zn > 0 ? (unsigned long long)p.zn : 1ull;  // no slices: every point is rejected
    const unsigned long long limit = (unsigned long long)batch * (unsigned long long)(p.xn * (long long)p.yn) * zn;
    return (limit > 0 && limit < 0xFFFFFFFFull) ? (unsigned)limit : 0u;
}

static int sort_end_bit(unsigned invalid_key)
{
    int bits = 0;
    while (bits < 32 && (invalid_key >> bits) != 0u) bits++;
    return bits < 1 ? 1 : bits;
}

size_t sort_temp_bytes(size_t n, unsigned invalid_key)
{
    size_t bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const unsigned *)nullptr, (unsigned *)nullptr, (const unsigned *)nullptr,
                                    (unsigned *)nullptr, (long long)n, 0, sort_end_bit(invalid_key));
    return bytes;
}

int launch_sort_keys(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, int batch, int max_points,
                     unsigned *keys, unsigned *vals, unsigned *fv_winner, unsigned invalid_key, cudaStream_t s)
{
    int b = (max_points + kSortThreads - 1) / kSortThreads;
    b = b < 1 ? 1 : (b > 4096 ? 4096 : b);
    sort_keys_kernel<<<dim3(b, batch), kSortThreads, 0, s>>>(pts, offs, p, keys, vals, fv_winner, invalid_key);
    return 1;
}

int launch_sort_pairs(void *temp, size_t temp_bytes, const unsigned *keys_in, unsigned *keys_out, const unsigned *vals_in,
                      unsigned *vals_out, size_t n, unsigned invalid_key, cudaStream_t s)
{
    if (n == 0) return 0;
    cub::DeviceRadixSort::SortPairs(temp, temp_bytes, keys_in, keys_out, vals_in, vals_out, (long long)n, 0,
                                    sort_end_bit(invalid_key), s);
    return 0;  // library kernels: timed, not counted as launches of this library
}

int launch_sort_reduce(const float4 *pts, const FrameOffsets &offs, const ViewParams &p, const float *lut,
                       const unsigned *skeys, const unsigned *svals, size_t n, unsigned invalid_key, float *top,
                       cudaStream_t s)
{
    if (n == 0) return 0;
    const size_t blocks = (n + kSortThreads - 1) / kSortThreads;
    sort_reduce_kernel<<<(unsigned)blocks, kSortThreads, 0, s>>>(pts, offs, p, lut, skeys, svals, n, invalid_key, top);
    return 1;
}

MV3D_SELF_CHECK_READER(selfcheck_sort)

} // namespace mv3d
