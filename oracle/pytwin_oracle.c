/*
 * pytwin_oracle.c -- C helpers of the Python-twin oracle (TEST INFRASTRUCTURE ONLY, see oracle/pytwin.py).
 *
 * The reference's Python encoders (src/data.py:113-180 lidar_to_front_fast, :296-367 lidar_to_top) run on float32
 * numpy arrays. Their only transcendental on the data path is np.arctan2 on float32, which the numpy of the
 * reference's time (1.13) forwarded to the C library's atan2f. glibc's atan2f (up to 2.40, incl. this image's 2.39)
 * is the fdlibm single-precision routine -- not correctly rounded, so "atan2 in double, rounded" differs from it in
 * ~20 % of the results and a newer numpy's AVX-512 loop differs again. The CUDA path therefore carries a port of
 * that routine; this file holds
 *   - libm_atan2f_array : the C library's own atan2f (the thing the oracle is defined by), and
 *   - fdlibm_atan2f_array: the same restatement the CUDA path uses, so tests can pin port == libm on the CPU.
 * Algorithm: FreeBSD/fdlibm e_atan2f.c + s_atanf.c (Sun Microsystems; "Permission to use, copy, modify, and
 * distribute this software is freely granted, provided that this notice is preserved"), restated here.
 *
 * Build: gcc -O2 -ffp-contract=off -shared -fPIC pytwin_oracle.c -o _build/libpytwin_oracle.so -lm
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

static inline uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static inline float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

static const float atanhi[4] = {4.6364760399e-01f, 7.8539812565e-01f, 9.8279368877e-01f, 1.5707962513e+00f};
static const float atanlo[4] = {5.0121582440e-09f, 3.7748947079e-08f, 3.4473217170e-08f, 7.5497894159e-08f};
static const float aT[11] = {3.3333334327e-01f, -2.0000000298e-01f, 1.4285714924e-01f, -1.1111110449e-01f,
                             9.0908870101e-02f, -7.6918758452e-02f, 6.6610731184e-02f, -5.8335702866e-02f,
                             4.9768779427e-02f, -3.6531571299e-02f, 1.6285819933e-02f};

static float fd_atanf(float x)
{
    const int32_t hx = (int32_t)f2u(x);
    const int32_t ix = hx & 0x7fffffff;
    int id;
    if (ix >= 0x4c000000) { /* |x| >= 2^25 (glibc; classic fdlibm has 2^34) */
        if (ix > 0x7f800000) return x + x;
        return hx > 0 ? atanhi[3] + atanlo[3] : -atanhi[3] - atanlo[3];
    }
    if (ix < 0x3ee00000) { /* |x| < 0.4375 */
        if (ix < 0x31000000) return x; /* |x| < 2^-29 */
        id = -1;
    } else {
        x = fabsf(x);
        if (ix < 0x3f980000) {     /* |x| < 1.1875 */
            if (ix < 0x3f300000) { id = 0; x = (2.0f * x - 1.0f) / (2.0f + x); }  /* 7/16 <= |x| < 11/16 */
            else { id = 1; x = (x - 1.0f) / (x + 1.0f); }                         /* 11/16 <= |x| < 19/16 */
        } else {
            if (ix < 0x401c0000) { id = 2; x = (x - 1.5f) / (1.0f + 1.5f * x); }  /* |x| < 2.4375 */
            else { id = 3; x = -1.0f / x; }
        }
    }
    const float z = x * x;
    const float w = z * z;
    const float s1 = z * (aT[0] + w * (aT[2] + w * (aT[4] + w * (aT[6] + w * (aT[8] + w * aT[10])))));
    const float s2 = w * (aT[1] + w * (aT[3] + w * (aT[5] + w * (aT[7] + w * aT[9]))));
    if (id < 0) return x - x * (s1 + s2);
    const float r = atanhi[id] - ((x * (s1 + s2) - atanlo[id]) - x);
    return hx < 0 ? -r : r;
}

static float fd_atan2f(float y, float x)
{
    static const float tiny = 1.0e-30f, pi_o_4 = 7.8539818525e-01f, pi_o_2 = 1.5707963705e+00f,
                       pi = 3.1415927410e+00f, pi_lo = -8.7422776573e-08f;
    const int32_t hx = (int32_t)f2u(x), hy = (int32_t)f2u(y);
    const int32_t ix = hx & 0x7fffffff, iy = hy & 0x7fffffff;
    if (ix > 0x7f800000 || iy > 0x7f800000) return x + y;
    if (hx == 0x3f800000) return fd_atanf(y);
    const int m = ((hy >> 31) & 1) | ((hx >> 30) & 2);
    if (iy == 0) {
        switch (m) {
        case 0: case 1: return y;
        case 2: return pi + tiny;
        default: return -pi - tiny;
        }
    }
    if (ix == 0) return hy < 0 ? -pi_o_2 - tiny : pi_o_2 + tiny;
    if (ix == 0x7f800000) {
        if (iy == 0x7f800000) {
            switch (m) {
            case 0: return pi_o_4 + tiny;
            case 1: return -pi_o_4 - tiny;
            case 2: return 3.0f * pi_o_4 + tiny;
            default: return -3.0f * pi_o_4 - tiny;
            }
        }
        switch (m) {
        case 0: return 0.0f;
        case 1: return -0.0f;
        case 2: return pi + tiny;
        default: return -pi - tiny;
        }
    }
    if (iy == 0x7f800000) return hy < 0 ? -pi_o_2 - tiny : pi_o_2 + tiny;
    const int32_t k = (iy - ix) >> 23;
    float z;
    if (k > 60) z = pi_o_2 + 0.5f * pi_lo;
    else if (hx < 0 && k < -60) z = 0.0f;
    else z = fd_atanf(fabsf(y / x));
    switch (m) {
    case 0: return z;
    case 1: return u2f(f2u(z) ^ 0x80000000u);
    case 2: return pi - (z - pi_lo);
    default: return (z - pi_lo) - pi;
    }
}

void fdlibm_atan2f_array(const float *y, const float *x, float *out, long n)
{
    for (long i = 0; i < n; i++) out[i] = fd_atan2f(y[i], x[i]);
}

void libm_atan2f_array(const float *y, const float *x, float *out, long n)
{
    for (long i = 0; i < n; i++) out[i] = atan2f(y[i], x[i]);
}

/* numpy's float32 floor_divide (npy_divmodf, numpy/core/src/npymath/npy_math_internal.h.src), the `//` of
 * src/data.py:321-322: not floorf(a / b) -- the quotient is rebuilt from fmodf. Restated for the CUDA port's test. */
float npy_floor_dividef(float a, float b)
{
    if (b == 0.0f) return a / b;
    float mod = fmodf(a, b);
    float div = (a - mod) / b;
    if (mod != 0.0f) {
        if ((b < 0.0f) != (mod < 0.0f)) div -= 1.0f;
    }
    if (div != 0.0f) {
        float fl = floorf(div);
        if (div - fl > 0.5f) fl += 1.0f;
        return fl;
    }
    return copysignf(0.0f, a / b);
}

void npy_floor_dividef_array(const float *a, float b, float *out, long n)
{
    for (long i = 0; i < n; i++) out[i] = npy_floor_dividef(a[i], b);
}
