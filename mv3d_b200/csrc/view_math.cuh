// view_math.cuh -- bit-exact device arithmetic of the MV3D view encoder.
//
// Every helper reproduces one expression of the reference C routine with the same IEEE operations in
// the same order (x86-64 SSE2 code: one rounding per float/double op, no FMA contraction, cvttss2si /
// cvttsd2si for the (int) casts). Explicit _rn intrinsics are used so nvcc never fuses or reorders.
// Reference lines: LidarTopPreprocess.c (T) and LidarTopAndFrontPreprocess.c (TF) under
// src/lidar_data_preprocess/Python_to_C_Interface/ver3/.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mv3d {

// Everything the kernels need, by value in the kernel parameter space.
struct ViewParams {
    // bird's-eye view
    float x_min, y_min, z_min;
    float x_div, y_div, z_div;
    int xn, yn, zn, planes;       // planes = zn + 2
    int quant_floor;              // 0: (int)q [T:107-109]   1: (int)floor(q) [TF:124-126]
    int want_top, want_front;
    long long top_frame_floats;   // xn*yn*planes
    // output tiling (tile strategy): a tile is tile_cells consecutive output cells = one contiguous chunk
    int tile_cells, tiles_per_frame;
    // front view
    float z_span;                 // z_MAX - z_MIN, rounded to float [TF:185]
    float rad_to_deg;             // (float)(180/3.141596) [TF:36]
    float phi_div, theta_div;
    float phi_scale, theta_scale; // rad_to_deg / div in float: only for the approximate fast path below
    float r_size_f, c_size_f;     // the float bounds of TF:176
    int rn, cn, fn, center_r, center_c;
    int front_pixels;             // rn*cn
    long long front_frame_floats; // rn*cn*fn
};

constexpr int kMaxBatchPerLaunch = 64;
struct FrameOffsets {            // first point of each sweep in the packed point array
    int off[kMaxBatchPerLaunch + 1];
};
constexpr int kDensityLutSize = 32;  // min(1, (float)(logf(c+1)/log(32))) for c = 0..31, built on the host with libm

constexpr uint32_t kInvalidKey = 0xFFFFFFFFu;
constexpr uint32_t kNoCell = 0xFFFFFFFFu;   // unused slot of the compact read-back list
constexpr uint32_t kRawTag = 0x80000000u;   // marks an un-finalised height slot in the atomic strategy

// x86 cvttss2si: NaN / out-of-range -> INT_MIN (CUDA's own conversion saturates and maps NaN to 0).
__device__ __forceinline__ int cvtt_f32(float v)
{
    return (v >= -2147483648.0f && v < 2147483648.0f) ? __float2int_rz(v) : INT32_MIN;
}
// x86 cvttsd2si.
__device__ __forceinline__ int cvtt_f64(double v)
{
    return (v > -2147483649.0 && v < 2147483648.0) ? __double2int_rz(v) : INT32_MIN;
}

// Cell coordinate along one axis: two separately rounded float ops, then truncate or floor.
__device__ __forceinline__ int quantise(float p, float lo, float div, int quant_floor)
{
    float q = __fdiv_rn(__fsub_rn(p, lo), div);
    if (quant_floor) q = floorf(q);
    return cvtt_f32(q);
}

struct CellHit {
    bool keep;       // passed the gate of T:112 / TF:129
    uint32_t cell;   // output cell (X-1)*yn + (Y-1)
    int slice;       // Z
    float h;         // z - z_MIN, the comparison key of T:115,141 / TF:133,159
};

__device__ __forceinline__ CellHit classify(const float4 &pt, const ViewParams &p)
{
    CellHit c;
    const int X = quantise(pt.x, p.x_min, p.x_div, p.quant_floor);
    const int Y = quantise(pt.y, p.y_min, p.y_div, p.quant_floor);
    const int Z = quantise(pt.z, p.z_min, p.z_div, p.quant_floor);
    c.keep = (X >= 1 && Y >= 1 && Z >= 0 && X < p.xn && Y < p.yn && Z < p.zn);
    c.cell = c.keep ? (uint32_t)((X - 1) * p.yn + (Y - 1)) : kInvalidKey;
    c.slice = Z;
    c.h = __fsub_rn(pt.z, p.z_min);
    return c;
}

// Height-slice output value: h / z_DIV - k  [T:205; TF:143 then :258].
__device__ __forceinline__ float slice_value(float h, float z_div, int k)
{
    return __fsub_rn(__fdiv_rn(h, z_div), (float)k);
}

// Ordering key for "highest point, first in input order": max over (h bits, ~index). h > 0 here, so the
// float's bit pattern orders like an unsigned integer.
__device__ __forceinline__ unsigned long long cell_key(float h, uint32_t idx)
{
    return ((unsigned long long)__float_as_uint(h) << 32) | (unsigned long long)(0xFFFFFFFFu - idx);
}
__device__ __forceinline__ uint32_t cell_key_index(unsigned long long key)
{
    return 0xFFFFFFFFu - (uint32_t)(key & 0xFFFFFFFFull);
}

// Angular bin (int)(atan2(num, den) * RAD_TO_DEG / DIV), the reference's double expression [TF:173-174].
__device__ __forceinline__ int angle_bin_exact(double num, double den, float rad_to_deg, float div)
{
    return cvtt_f64(__ddiv_rn(__dmul_rn(atan2(num, den), (double)rad_to_deg), (double)div));
}

// Same bin from single-precision arithmetic, when that is provably enough. The float quotient qf differs from
// the reference's double value by at most |q|*4e-7 + |scale|*0.6e-7 (2-ulp atan2f, <=1.25-ulp sqrt/arguments,
// two roundings for the scale); the bin is accepted only if qf is at least eps = |qf|*4e-6 + |scale|*1e-6
// (about 10x that bound) away from every integer, where truncation cannot differ. Otherwise the caller takes
// the exact double path. About 0.1 % of points do; the result is bit-identical either way.
__device__ __forceinline__ bool angle_bin_fast(float num, float den, float scale, int &bin)
{
    const float qf = __fmul_rn(atan2f(num, den), scale);
    const float eps = fabsf(qf) * 4e-6f + fabsf(scale) * 1e-6f;
    if (!(fabsf(qf) < 1.0e6f) || fabsf(qf - rintf(qf)) <= eps) return false;
    bin = __float2int_rz(qf);
    return true;
}

// Front-view pixel (linear index r_coord*C + c_coord) of a point that passed the BEV gate, or -1 when the
// reference would not store it in bounds. The reference's inclusive upper bounds are kept, so c_coord == C
// aliases pixel (r_coord+1, 0) exactly as its linear index does, and anything at or beyond R*C (the
// reference's out-of-bounds store) is dropped.
__device__ __forceinline__ int front_pixel(const float4 &pt, const ViewParams &p)
{
    int R, C;
    const float planar_f = sqrtf(__fadd_rn(__fmul_rn(pt.x, pt.x), __fmul_rn(pt.y, pt.y)));
    if (!angle_bin_fast(pt.z, planar_f, p.phi_scale, R)) {
        const double dx = (double)pt.x, dy = (double)pt.y;
        const double planar = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
        R = angle_bin_exact((double)pt.z, planar, p.rad_to_deg, p.phi_div);
    }
    if (!angle_bin_fast(pt.y, pt.x, p.theta_scale, C))
        C = angle_bin_exact((double)pt.y, (double)pt.x, p.rad_to_deg, p.theta_div);
    const long long r = (long long)R + p.center_r, c = (long long)C + p.center_c;
    if (!((float)r <= p.r_size_f && (float)c <= p.c_size_f && r >= 0 && c >= 0)) return -1;
    const long long pix = r * p.cn + c;
    return (pix < (long long)p.front_pixels) ? (int)pix : -1;
}

// The all-double evaluation of the same pixel (no fast path); used by the debug entry point to cross-check.
__device__ __forceinline__ int front_pixel_exact(const float4 &pt, const ViewParams &p)
{
    const double dx = (double)pt.x, dy = (double)pt.y;
    const double planar = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
    const int R = angle_bin_exact((double)pt.z, planar, p.rad_to_deg, p.phi_div);
    const int C = angle_bin_exact(dy, dx, p.rad_to_deg, p.theta_div);
    const long long r = (long long)R + p.center_r, c = (long long)C + p.center_c;
    if (!((float)r <= p.r_size_f && (float)c <= p.c_size_f && r >= 0 && c >= 0)) return -1;
    const long long pix = r * p.cn + c;
    return (pix < (long long)p.front_pixels) ? (int)pix : -1;
}

// The three front-view channel values of a point [TF:185,188,192], as the floats the reference stores.
__device__ __forceinline__ void front_values(const float4 &pt, const ViewParams &p, float &v0, float &v1, float &v2)
{
    const double dx = (double)pt.x, dy = (double)pt.y;
    const double planar = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
    v0 = (float)cvtt_f32(__fmul_rn(__fdiv_rn(__fsub_rn(pt.z, p.z_min), p.z_span), 255.0f));
    v1 = (float)cvtt_f64(__dmul_rn(__ddiv_rn(planar, 120.0), 255.0));
    v2 = (float)cvtt_f32(__fmul_rn(__fsub_rn(pt.w, 0.0f), 255.0f));
}

// Streaming 128-bit point load: every point is read once per pass, keep it out of L1.
__device__ __forceinline__ float4 load_point(const float4 *p)
{
    float4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

} // namespace mv3d
