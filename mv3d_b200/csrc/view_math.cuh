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
    float x_inv, y_inv, z_inv;    // 1/div rounded to float (NaN when div is not a positive normal number): fast path of quantise()
    int xn, yn, zn, planes;       // planes = zn + 2
    int quant_floor;              // 0: (int)q [T:107-109]   1: (int)floor(q) [TF:124-126]
    int py;                       // 1: semantics of the reference's PYTHON lidar_to_top (src/data.py:296-367), see py_classify
    int center_car;               // drop points with |x| <= 2.35 and |y| <= 1.05 first (filter_center_car, src/data.py:224-227,266-267,313-314)
    int flip_xy;                  // C routine only: store cell (X-1, Y-1) at [xn-X][yn-Y], the np.flipud(np.fliplr(.)) of src/data.py:290
    float x_max, y_max, z_max;    // the Python encoders crop with open intervals on all three axes
    int want_top, want_front;
    long long top_frame_floats;   // xn*yn*planes
    // output tiling (tile strategy): a tile is tile_cells consecutive output cells = one contiguous chunk
    int tile_cells, tiles_per_frame;
    unsigned long long tile_magic; // ceil(2^48 / tile_cells) << 16: cell / tile_cells == umul64hi(cell, tile_magic), see tile_of()
    // front view
    float z_span;                 // z_MAX - z_MIN, rounded to float [TF:185]
    float rad_to_deg;             // (float)(180/3.141596) [TF:36]
    float phi_div, theta_div;
    float phi_scale, theta_scale; // rad_to_deg / div in float: only for the approximate fast path below
    float phi_eps, theta_eps;     // |scale| * 2.5e-6: the part of the fast path's error bound that does not grow with the bin
    float r_size_f, c_size_f;     // the float bounds of TF:176
    int rn, cn, fn, center_r, center_c;
    int front_pixels;             // rn*cn
    long long front_frame_floats; // rn*cn*fn
};

// Tile of an output cell without an integer division: for cell < 2^31 and tile_cells < 2^16,
// floor(cell * ceil(2^48 / d) / 2^48) == cell / d (the error term cell * d / 2^48 stays below 1 / d).
__host__ __device__ inline unsigned long long tile_magic_for(int tile_cells)
{
    const unsigned long long d = (unsigned long long)(tile_cells > 0 ? tile_cells : 1);
    return (((1ull << 48) + d - 1) / d) << 16;
}
__device__ __forceinline__ uint32_t tile_of(uint32_t cell, const ViewParams &p)
{
    return (uint32_t)__umul64hi((unsigned long long)cell, p.tile_magic);
}

constexpr int kMaxBatchPerLaunch = 64;
struct FrameOffsets {            // first point of each sweep in the packed point array
    int off[kMaxBatchPerLaunch + 1];
};
constexpr int kDensityLutSize = 32;  // min(1, (float)(logf(c+1)/log(32))) for c = 0..31, built on the host with libm

constexpr uint32_t kInvalidKey = 0xFFFFFFFFu;
constexpr uint32_t kLowerFlag = 0x8000u;     // Python twin, slice field of a binned record: also closes the slice below
constexpr uint32_t kNoSlice = 0x7FFFu;       // Python twin: floor(qz) is past the last slice
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

// Cell coordinate along one axis: two separately rounded float ops, then truncate or floor [T:107-109 / TF:124-126].
// The exact form (an IEEE division) is what the reference computes.
__device__ __forceinline__ int quantise_exact(float a, float div, int quant_floor)
{
    float q = __fdiv_rn(a, div);
    if (quant_floor) q = floorf(q);
    return cvtt_f32(q);
}
// Same integer from a multiplication by the rounded reciprocal, when that is provably enough: qf = RN(a * RN(1/div))
// differs from the reference's RN(a / div) by at most |q| * 3 * 2^-24 = |q| * 1.8e-7 (three roundings). Truncation
// and floor can only differ from the reference's when an integer lies between the two values, so qf is accepted when
// it is at least |qf| * 1e-6 (5x that bound) away from every integer; otherwise -- on cell boundaries, about one point
// in a thousand on real sweeps -- the IEEE division decides. NaN / huge / denormal-reciprocal cases take the exact
// path too (inv is NaN when div is not a positive normal float). Bit-identical either way; the per-point tests
// compare 20 M random points and the boundary-snapped fuzz clouds with the oracle.
__device__ __forceinline__ int quantise(float p, float lo, float div, float inv, int quant_floor)
{
    const float a = __fsub_rn(p, lo);
    const float qf = __fmul_rn(a, inv);
    const float aq = fabsf(qf);
    if (!(aq < 4.0e6f) || fabsf(__fsub_rn(qf, rintf(qf))) <= __fmul_rn(aq, 1.0e-6f)) return quantise_exact(a, div, quant_floor);
    return quant_floor ? __float2int_rd(qf) : __float2int_rz(qf);
}

struct CellHit {
    bool keep;       // passed the gate of T:112 / TF:129
    uint32_t cell;   // output cell (X-1)*yn + (Y-1)
    int slice;       // Z
    float h;         // z - z_MIN, the comparison key of T:115,141 / TF:133,159
    bool lower;      // Python twin only: the point also closes slice Z-1 (its qz is integral)
};

// ---- the reference's PYTHON bird's-eye view (src/data.py:296-367), on float32 like numpy computes it ----

// numpy's float32 floor_divide (npy_divmodf): the quotient is rebuilt from fmodf, it is NOT floorf(a / b).
// fmodf is exact in CUDA as in C. Restated on the CPU in oracle/pytwin_oracle.c and checked against numpy.
__device__ __forceinline__ float npy_floor_dividef(float a, float b)
{
    if (b == 0.0f) return __fdiv_rn(a, b);
    const float mod = fmodf(a, b);
    float div = __fdiv_rn(__fsub_rn(a, mod), b);
    if (mod != 0.0f && ((b < 0.0f) != (mod < 0.0f))) div = __fsub_rn(div, 1.0f);
    if (div != 0.0f) {
        float fl = floorf(div);
        if (__fsub_rn(div, fl) > 0.5f) fl = __fadd_rn(fl, 1.0f);
        return fl;
    }
    return copysignf(0.0f, __fdiv_rn(a, b));
}

// Open-interval crop on all three axes (src/data.py:298-311 = :130-143); NaN fails every comparison.
__device__ __forceinline__ bool py_crop(const float4 &pt, const ViewParams &p)
{
    return pt.x > p.x_min && pt.x < p.x_max && pt.y > p.y_min && pt.y < p.y_max && pt.z > p.z_min && pt.z < p.z_max;
}


// Gate, cell and slice key of the Python encoder:
//   crop (open intervals) [+ centre-car filter], qx/qy by numpy floor_divide, qz = (z - z_MIN) / z_DIV kept fractional
//   (:321-324); only 0 <= qx < Xn, 0 <= qy < Yn are ever visited (:341,:347); the cell is written flipped at
//   top[-qx-1, -qy-1] (:345,:352). The comparison key is qz (> 0 after the crop): slice z takes every point with
//   z <= qz <= z + 1, BOTH ends inclusive (:359), so an integral qz also closes the slice below. qz bits order
//   like unsigned integers, exactly as h does for the C routine, so the same reduction serves both.
__device__ __forceinline__ CellHit py_classify(const float4 &pt, const ViewParams &p)
{
    CellHit c;
    c.keep = false;
    c.cell = kInvalidKey;
    c.slice = 0;
    c.h = 0.0f;
    c.lower = false;
    if (!py_crop(pt, p)) return c;
    if (p.center_car && !(fabsf(pt.x) > 2.35f || fabsf(pt.y) > 1.05f)) return c;  // f32(4.7/2), f32(2.1/2)
    const int qx = cvtt_f32(npy_floor_dividef(__fsub_rn(pt.x, p.x_min), p.x_div));
    const int qy = cvtt_f32(npy_floor_dividef(__fsub_rn(pt.y, p.y_min), p.y_div));
    if (!(qx >= 0 && qx < p.xn && qy >= 0 && qy < p.yn)) return c;
    const float qz = __fdiv_rn(__fsub_rn(pt.z, p.z_min), p.z_div);
    const float fl = floorf(qz);
    c.keep = true;
    c.cell = (uint32_t)((p.xn - 1 - qx) * p.yn + (p.yn - 1 - qy));
    c.slice = cvtt_f32(fl);
    c.h = qz;
    c.lower = (qz == fl) && c.slice >= 1;
    return c;
}

// kPy selects the Python encoder's semantics at compile time: the kernels of the tile strategy are instantiated for
// both, so the C-routine path carries none of the other's branches or registers.
template <bool kPy>
__device__ __forceinline__ CellHit classify_as(const float4 &pt, const ViewParams &p)
{
    if (kPy) return py_classify(pt, p);
    CellHit c;
    c.lower = false;
    if (p.center_car && !(fabsf(pt.x) > 2.35f || fabsf(pt.y) > 1.05f)) {  // removed before the C routine is called
        c.keep = false;
        c.cell = kInvalidKey;
        c.slice = 0;
        c.h = 0.0f;
        return c;
    }
    const int X = quantise(pt.x, p.x_min, p.x_div, p.x_inv, p.quant_floor);
    const int Y = quantise(pt.y, p.y_min, p.y_div, p.y_inv, p.quant_floor);
    const int Z = quantise(pt.z, p.z_min, p.z_div, p.z_inv, p.quant_floor);
    c.keep = (X >= 1 && Y >= 1 && Z >= 0 && X < p.xn && Y < p.yn && Z < p.zn);
    c.cell = !c.keep ? kInvalidKey
                     : p.flip_xy ? (uint32_t)((p.xn - X) * p.yn + (p.yn - Y)) : (uint32_t)((X - 1) * p.yn + (Y - 1));
    c.slice = Z;
    c.h = __fsub_rn(pt.z, p.z_min);
    return c;
}

__device__ __forceinline__ CellHit classify(const float4 &pt, const ViewParams &p)
{
    return p.py ? classify_as<true>(pt, p) : classify_as<false>(pt, p);
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

// ---- front-view bins [TF:173-174] ---------------------------------------------------------------------------------
// The reference file is C++ (<math.h> = libstdc++'s wrapper, `using namespace std`), so the two bins have different
// types:
//   R = (int)(atan2(z, sqrt(pow(x,2)+pow(y,2))) * RAD_TO_DEG / phi_DIVISION)   pow(float,int) -> double: all double
//   C = (int)(atan2(y, x) * RAD_TO_DEG / theta_DIVISION)                       atan2(float,float) = atan2f: all float
// (`nm -D` on the reference build imports atan2 and atan2f; oracle/ref_probe.c checks both bins one point per call.)
__device__ __forceinline__ float fd_atan2f(float y, float x);

// Row bin, exact: the double expression, cvttsd2si.
__device__ __forceinline__ int row_bin_exact(double num, double den, float rad_to_deg, float div)
{
    return cvtt_f64(__ddiv_rn(__dmul_rn(atan2(num, den), (double)rad_to_deg), (double)div));
}
// Column bin, exact: the float expression on the C library's atan2f (fd_atan2f below restates glibc's, operation for
// operation), cvttss2si.
__device__ __forceinline__ int col_bin_exact(float y, float x, float rad_to_deg, float div)
{
    return cvtt_f32(__fdiv_rn(__fmul_rn(fd_atan2f(y, x), rad_to_deg), div));
}

// atan2 to within 4.2e-7 rad, ~25 instructions: t = min/max by the hardware reciprocal (2 ulp), atan(t) on [0, 1] by
// the 8-term odd polynomial of Abramowitz & Stegun 4.4.49 (|error| <= 2e-8; 9.3e-8 evaluated in float32 with FMAs),
// then the octant fix-ups (each one rounding of a value below pi, constants off by < 9e-8). Only ever used where an
// error bound decides whether its result is good enough (angle_bin_fast); never feeds a stored value.
__device__ __forceinline__ float approx_atan2f(float y, float x)
{
    const float ax = fabsf(x), ay = fabsf(y);
    const float mx = fmaxf(ax, ay), mn = fminf(ax, ay);
    const float t = __fdividef(mn, mx);  // 0/0 -> NaN: the caller's range test sends it to the exact path
    const float s = t * t;
    float r = 0.0028662257f;
    r = fmaf(r, s, -0.0161657367f);
    r = fmaf(r, s, 0.0429096138f);
    r = fmaf(r, s, -0.0752896400f);
    r = fmaf(r, s, 0.1065626393f);
    r = fmaf(r, s, -0.1420889944f);
    r = fmaf(r, s, 0.1999355085f);
    r = fmaf(r, s, -0.3333314528f);
    r = fmaf(r * s, t, t);
    if (ay > ax) r = 1.57079637f - r;
    if (x < 0.0f) r = 3.14159274f - r;
    return copysignf(r, y);
}

// A bin from single-precision arithmetic, when that is provably enough. qf = approx_atan2f * scale differs from the
// value the reference truncates by at most |scale| * 5.1e-7 (the angle, incl. the approximate planar range of the row
// bin) + |q| * 3.6e-7 (rounding of scale and product:
// 1.2e-7; the reference's own float roundings for the column bin: 2.4e-7; none for the double row bin). The bin is
// accepted only if qf is at least eps = |scale| * 2.5e-6 + |qf| * 2e-6 (5x / 5.5x those terms) away from every
// integer, where truncation cannot differ. Otherwise the caller takes the exact path: a few points in a thousand do,
// and the result is bit-identical either way (tests: 20 M random points, fast == exact == oracle).
__device__ __forceinline__ bool angle_bin_fast(float num, float den, float scale, float scale_eps, int &bin)
{
    const float qf = __fmul_rn(approx_atan2f(num, den), scale);
    const float aq = fabsf(qf);
    if (!(aq < 1.0e6f) || fabsf(qf - rintf(qf)) <= fmaf(aq, 2.0e-6f, scale_eps)) return false;
    bin = __float2int_rz(qf);
    return true;
}

// Front-view pixel (linear index r_coord*C + c_coord) of a point that passed the BEV gate, or -1 when the
// reference would not store it in bounds. The reference's inclusive upper bounds are kept, so c_coord == C
// aliases pixel (r_coord+1, 0) exactly as its linear index does, and anything at or beyond R*C (the
// reference's out-of-bounds store) is dropped.
__device__ __forceinline__ int front_pixel_of(int R, int C, const ViewParams &p)
{
    const long long r = (long long)R + p.center_r, c = (long long)C + p.center_c;
    if (!((float)r <= p.r_size_f && (float)c <= p.c_size_f && r >= 0 && c >= 0)) return -1;
    const long long pix = r * p.cn + c;
    return (pix < (long long)p.front_pixels) ? (int)pix : -1;
}
// The exact bins are the RARE path (a few points in a thousand): kept out of line, so that the double-precision atan2
// and the fdlibm atan2f do not take registers from the loops that call them.
static __device__ __noinline__ int row_bin_slow(float x, float y, float z, float rad_to_deg, float phi_div)
{
    const double dx = (double)x, dy = (double)y;
    const double planar = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
    return row_bin_exact((double)z, planar, rad_to_deg, phi_div);
}
static __device__ __noinline__ int col_bin_slow(float y, float x, float rad_to_deg, float theta_div)
{
    return col_bin_exact(y, x, rad_to_deg, theta_div);
}
__device__ __forceinline__ int row_bin_of(const float4 &pt, const ViewParams &p)
{
    return row_bin_slow(pt.x, pt.y, pt.z, p.rad_to_deg, p.phi_div);
}
__device__ __forceinline__ int front_pixel(const float4 &pt, const ViewParams &p)
{
    int R, C;
    float planar_f;  // approximate on purpose (1.8e-7 relative, 9e-8 rad on the angle): fast path only
    asm("sqrt.approx.f32 %0, %1;" : "=f"(planar_f) : "f"(fmaf(pt.x, pt.x, pt.y * pt.y)));
    if (!angle_bin_fast(pt.z, planar_f, p.phi_scale, p.phi_eps, R)) R = row_bin_of(pt, p);
    if (!angle_bin_fast(pt.y, pt.x, p.theta_scale, p.theta_eps, C)) C = col_bin_slow(pt.y, pt.x, p.rad_to_deg, p.theta_div);
    return front_pixel_of(R, C, p);
}

// The reference's own expressions for the same pixel (no fast path); used by the debug entry point to cross-check.
__device__ __forceinline__ int front_pixel_exact(const float4 &pt, const ViewParams &p)
{
    return front_pixel_of(row_bin_of(pt, p), col_bin_exact(pt.y, pt.x, p.rad_to_deg, p.theta_div), p);
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

// ---- the C library's atan2f ---------------------------------------------------------------------------------------
// Used by the column bin of the C routine's front view (above) and by the reference's PYTHON front view
// (src/data.py:113-180), where np.arctan2 on float32 was the C library's atan2f when the reference was written.
// glibc's (<= 2.40) is the fdlibm single-precision routine, which is not correctly rounded. This is that routine,
// operation for operation (e_atan2f.c / s_atanf.c; restated on the CPU in oracle/pytwin_oracle.c and pinned there to
// libm's atan2f).
__device__ __forceinline__ float fd_atanf(float x)
{
    const float atanhi[4] = {4.6364760399e-01f, 7.8539812565e-01f, 9.8279368877e-01f, 1.5707962513e+00f};
    const float atanlo[4] = {5.0121582440e-09f, 3.7748947079e-08f, 3.4473217170e-08f, 7.5497894159e-08f};
    const int32_t hx = __float_as_int(x);
    const int32_t ix = hx & 0x7fffffff;
    int id;
    if (ix >= 0x4c000000) {  // |x| >= 2^25
        if (ix > 0x7f800000) return __fadd_rn(x, x);
        const float r = __fadd_rn(atanhi[3], atanlo[3]);
        return hx > 0 ? r : -r;
    }
    if (ix < 0x3ee00000) {  // |x| < 0.4375
        if (ix < 0x31000000) return x;
        id = -1;
    } else {
        x = fabsf(x);
        if (ix < 0x3f980000) {
            if (ix < 0x3f300000) { id = 0; x = __fdiv_rn(__fsub_rn(__fmul_rn(2.0f, x), 1.0f), __fadd_rn(2.0f, x)); }
            else { id = 1; x = __fdiv_rn(__fsub_rn(x, 1.0f), __fadd_rn(x, 1.0f)); }
        } else {
            if (ix < 0x401c0000) { id = 2; x = __fdiv_rn(__fsub_rn(x, 1.5f), __fadd_rn(1.0f, __fmul_rn(1.5f, x))); }
            else { id = 3; x = __fdiv_rn(-1.0f, x); }
        }
    }
    const float z = __fmul_rn(x, x);
    const float w = __fmul_rn(z, z);
#define MV3D_H(c, t) __fadd_rn(c, __fmul_rn(w, t))
    const float s1 = __fmul_rn(z, MV3D_H(3.3333334327e-01f, MV3D_H(1.4285714924e-01f, MV3D_H(9.0908870101e-02f,
                     MV3D_H(6.6610731184e-02f, MV3D_H(4.9768779427e-02f, 1.6285819933e-02f))))));
    const float s2 = __fmul_rn(w, MV3D_H(-2.0000000298e-01f, MV3D_H(-1.1111110449e-01f, MV3D_H(-7.6918758452e-02f,
                     MV3D_H(-5.8335702866e-02f, -3.6531571299e-02f)))));
#undef MV3D_H
    const float xs = __fmul_rn(x, __fadd_rn(s1, s2));
    if (id < 0) return __fsub_rn(x, xs);
    const float r = __fsub_rn(atanhi[id], __fsub_rn(__fsub_rn(xs, atanlo[id]), x));
    return hx < 0 ? -r : r;
}

__device__ __forceinline__ float fd_atan2f(float y, float x)
{
    const float tiny = 1.0e-30f, pi_o_4 = 7.8539818525e-01f, pi_o_2 = 1.5707963705e+00f, pi = 3.1415927410e+00f,
                pi_lo = -8.7422776573e-08f;
    const int32_t hx = __float_as_int(x), hy = __float_as_int(y);
    const int32_t ix = hx & 0x7fffffff, iy = hy & 0x7fffffff;
    if (ix > 0x7f800000 || iy > 0x7f800000) return __fadd_rn(x, y);
    if (hx == 0x3f800000) return fd_atanf(y);
    const int m = ((hy >> 31) & 1) | ((hx >> 30) & 2);
    if (iy == 0) return m < 2 ? y : (m == 2 ? __fadd_rn(pi, tiny) : __fsub_rn(-pi, tiny));
    if (ix == 0) return hy < 0 ? __fsub_rn(-pi_o_2, tiny) : __fadd_rn(pi_o_2, tiny);
    if (ix == 0x7f800000) {
        if (iy == 0x7f800000) {
            const float q3 = __fmul_rn(3.0f, pi_o_4);
            return m == 0 ? __fadd_rn(pi_o_4, tiny) : m == 1 ? __fsub_rn(-pi_o_4, tiny)
                 : m == 2 ? __fadd_rn(q3, tiny) : __fsub_rn(-q3, tiny);
        }
        return m == 0 ? 0.0f : m == 1 ? -0.0f : m == 2 ? __fadd_rn(pi, tiny) : __fsub_rn(-pi, tiny);
    }
    if (iy == 0x7f800000) return hy < 0 ? __fsub_rn(-pi_o_2, tiny) : __fadd_rn(pi_o_2, tiny);
    const int32_t k = (iy - ix) >> 23;
    float z;
    if (k > 60) z = __fadd_rn(pi_o_2, __fmul_rn(0.5f, pi_lo));
    else if (hx < 0 && k < -60) z = 0.0f;
    else z = fd_atanf(fabsf(__fdiv_rn(y, x)));
    switch (m) {
    case 0: return z;
    case 1: return -z;
    case 2: return __fsub_rn(pi, __fsub_rn(z, pi_lo));
    default: return __fsub_rn(__fsub_rn(z, pi_lo), pi);
    }
}

// Parameters of the Python front view (src/config.py:32-42), by value in the kernel parameter space.
struct PyFrontParams {
    float ang_res, vert_res, velodyne_height;   // narrowed to float32 like numpy does with a Python scalar
    int r_offset, c_offset, r_min, r_max, c_min, c_max, width, height;
    int pixels;                                  // width * height
};

// Pixel c * HEIGHT + r of a point that passed the crop, or -1 (src/data.py:121-126,145-166): float32 atan2 over the
// float32 resolution, truncated; open-interval window on the bins, offsets, then the image bounds.
__device__ __forceinline__ int py_front_pixel(const float4 &pt, const PyFrontParams &f)
{
    const int c = cvtt_f32(__fdiv_rn(fd_atan2f(pt.y, pt.x), f.ang_res));
    const float planar = __fsqrt_rn(__fadd_rn(__fmul_rn(pt.x, pt.x), __fmul_rn(pt.y, pt.y)));
    const int r = cvtt_f32(__fdiv_rn(fd_atan2f(pt.z, planar), f.vert_res));
    if (!(f.c_min < c && c < f.c_max && f.r_min < r && r < f.r_max)) return -1;
    const int cc = c + f.c_offset, rr = r + f.r_offset;
    if (!(cc >= 0 && cc < f.width && rr >= 0 && rr < f.height)) return -1;
    return cc * f.height + rr;
}

// The three channel values a point adds to its pixel (src/data.py:115-120): clip(z + 1.73, 0), the norm of
// (x, y, z, intensity) -- sic, the intensity is inside the norm --, the intensity. All float32.
__device__ __forceinline__ float3 py_front_values(const float4 &pt, const PyFrontParams &f)
{
    float3 v;
    v.x = fmaxf(__fadd_rn(pt.z, f.velodyne_height), 0.0f);
    v.y = __fsqrt_rn(__fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(pt.x, pt.x), __fmul_rn(pt.y, pt.y)), __fmul_rn(pt.z, pt.z)),
                               __fmul_rn(pt.w, pt.w)));
    v.z = pt.w;
    return v;
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
